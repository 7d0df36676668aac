// Layout constants shared by the deformation-MLP kernels (forward, input-gradient chain, weight gradient).
//
// Network (utils/time_utils.py:56-201): PE(xyz) 63 | t-feature (timenet output 30, or PE(t) 21) -> "pe"
// block of pe_dim = 93 / 84 columns padded to 96; 8 layers of width 256, layer 5 takes [pe | h4];
// heads 256 -> up to 13 outputs padded to 16.
//
// Arithmetic: tcgen05 kind::f16 with fp32 accumulation on SPLIT operands.  Every fp32 value x is carried as
// hi = fp16(x), lo = fp16(x - hi) (22 significant bits together); a product A*W is evaluated as
// Ahi*Whi + Alo*Whi + Ahi*Wlo (the lo*lo term is below fp32 resolution).  Measured on B200
// (profiles/r2_numerics_split.json): rms 3.4e-7 / max 5.9e-7 relative on a K=256 layer — the accuracy class of an
// fp32 cuBLAS GEMM (2.0e-7 / 5.3e-7) — against 2.7e-4 for single-pass TF32, at 1.5x the tensor time of TF32
// (kind::f16 runs at twice the kind::tf32 rate).  This is what makes the ReLU pattern — and with it the parameter
// gradients — agree with an fp32/fp64 evaluation (DESIGN.md "MLP precision").  Weights are pre-scaled by
// 2^kWScaleLog2 so that hi and lo stay in fp16's normal range; the epilogues undo the scale.
//
// One tile = 128 points.  All 16-bit matrix operands are panelised SWIZZLE_128B images (umma.cuh, img_off16).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"
#include "umma.cuh"

namespace d3g {

constexpr int kRows = 128;                 // points per tile
constexpr int kWScaleLog2 = 8;             // weights are packed as fp16(W * 2^8)
constexpr float kWScale = 256.0f, kWScaleInv = 1.0f / 256.0f;
constexpr uint32_t kPanelBytes = 16384;    // 128 rows x 128 B (64 fp16, or 32 hi | 32 lo)
constexpr uint32_t kPeSmemBytes = 3 * kPanelBytes;  // pe block as A operand: 3 k-chunks of [32 hi | 32 lo]
constexpr uint32_t kPeBytes = 2 * kPanelBytes;      // pe block saved for wgrad: 2 panels of 64 features (96 used)
constexpr uint32_t kHBytes = 4 * kPanelBytes;       // one 128x256 fp16 activation / gradient tile image
constexpr uint32_t kStageBytes = 32768;    // one weight stage: 256 rows x (32 k hi | 32 k lo)
constexpr int kFwdStages = 63;             // L0:3 L1-4:32 L5pe:3 L5h:8 L6-7:16 heads:1
constexpr int kBwdStages = 57;             // headsT:1 + 7 layers x 8
constexpr int kBiasFloats = 8 * 256 + 16;
constexpr size_t kFwdBytes = (size_t)62 * kStageBytes + 16384;
constexpr size_t kBiasOff = kFwdBytes;
constexpr size_t kBiasBytes = 8448;  // kBiasFloats*4 rounded up to 256; the tail holds the 9 backward gain exponents
constexpr size_t kGainOff = kBiasOff + kBiasFloats * 4;   // int32[9]: trunk layers 0..7 (1..7 used), [8] heads
constexpr size_t kBwdOff = kBiasOff + kBiasBytes;
constexpr size_t kBwdBytes = (size_t)kBwdStages * kStageBytes;
constexpr size_t kPackedBytes = kBwdOff + kBwdBytes;

// split an fp32 value pair into packed fp16 hi / lo words (low half = first element)
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(x0, x1);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
  hi = *(const uint32_t*)&h;
  lo = *(const uint32_t*)&l;
}
// Backward range management (fp16 holds 2^-14 .. 2^16 at full precision; gradients arrive at any magnitude and
// shrink or grow along the chain).  Two exact power-of-two factors keep the chain centred: gscale, from the largest
// |head cotangent| of the call (mlp_amax_kernel), and one gain per chain step, 2^e with e from the Frobenius norm of
// that step's weights (what a random-direction gradient is multiplied by; default init: 0.41 per layer -> e = 1),
// computed by pack_gains_kernel whenever the weights are re-packed.  The dZ image of layer l therefore carries
// 2^E_l, E_7 = s + e_heads, E_l = E_{l+1} + e_{l+1}; every consumer multiplies by 2^-E_l (exact) on the way out.
struct GradScales {
  float step_mul[8];   // epilogue multiplier of chain step j = 2^-kWScaleLog2 * 2^e
  float dz_inv[8];     // 2^-E_l for the dZ image of layer l
  float gs, gs_inv;    // scale of the head cotangents (dout image)
};
__device__ __forceinline__ float pow2i(int e) {
  e = e < -120 ? -120 : (e > 120 ? 120 : e);
  return __uint_as_float((uint32_t)(e + 127) << 23);
}
__device__ __forceinline__ GradScales grad_scales(uint32_t amax_bits, int target, const int* __restrict__ gain_e) {
  GradScales g;
  int s = 0;
  if (amax_bits != 0u && amax_bits < 0x7F800000u) s = target - ((int)(amax_bits >> 23) - 127);
  s = s < -100 ? -100 : (s > 100 ? 100 : s);
  g.gs = pow2i(s); g.gs_inv = pow2i(-s);
  int E = s + gain_e[8];
  g.step_mul[0] = pow2i(gain_e[8] - kWScaleLog2);
  g.dz_inv[7] = pow2i(-E);
#pragma unroll
  for (int j = 1; j < 8; j++) {     // step j forms dZ_{7-j} from dZ_{8-j} * W_{8-j}
    const int e = gain_e[8 - j];
    g.step_mul[j] = pow2i(e - kWScaleLog2);
    E += e;
    g.dz_inv[7 - j] = pow2i(-E);
  }
  return g;
}
__device__ __forceinline__ uint32_t pack2(float x0, float x1) {
  const __half2 h = __floats2half2_rn(x0, x1);
  return *(const uint32_t*)&h;
}

// forward stage -> (layer id, k panel).  layer ids: 0..7 trunk (5 = h part of the skip layer),
// 8 = pe part of layer 5, 9 = heads.
__host__ __device__ inline void fwd_stage_info(int st, int& layer, int& kpanel) {
  if (st < 3) { layer = 0; kpanel = st; }
  else if (st < 35) { layer = 1 + (st - 3) / 8; kpanel = (st - 3) % 8; }
  else if (st < 38) { layer = 8; kpanel = st - 35; }
  else if (st < 46) { layer = 5; kpanel = st - 38; }
  else if (st < 62) { layer = 6 + (st - 46) / 8; kpanel = (st - 46) % 8; }
  else { layer = 9; kpanel = 0; }
}
// backward stage order: headsT, then layers 7,6,5(h part),4,3,2,1 transposed
__host__ __device__ inline void bwd_stage_info(int st, int& layer, int& kpanel) {
  if (st == 0) { layer = 9; kpanel = 0; }
  else { layer = 7 - (st - 1) / 8; kpanel = (st - 1) % 8; }
}

// ---- workspace layout -------------------------------------------------------------------------
// Saved tiles are fp16 [point][feature] images (img_off16 with rows = points): exactly what the weight-gradient
// kernel feeds to the tensor core as MN-major operands with K = points, so backward never re-arranges them.
struct MlpWs {
  int32_t* status;     // [0] abort flag, [1] "rows carry different time values" (set on the device by forward)
  float* tfeat;        // [rows_t, 32]   time feature fed into pe columns 63..
  float* thidden;      // [rows_t, 256]  timenet hidden (post-ReLU), blender only
  uint8_t* pe_img;     // [tiles] x 32 KB            pe block (96 of 128 columns)
  uint8_t* h_img;      // [8][tiles] x 64 KB         post-ReLU activations h_0..h_7 (fp16 hi part)
  uint32_t* mask;      // [8][tiles][8 chunks][128 rows]  bit j = (h[row][chunk*32+j] > 0)
  float* heads_out;    // [tiles*128, 16] head pre-outputs (incl. bias)
  uint8_t* dz_img;     // [8][tiles] x 64 KB         gscale * dL/d(pre-activation) tiles (fp16), written by backward
  uint8_t* dout_img;   // [tiles] x 16 KB            gscale * head cotangents (one 64-column panel, 16 used)
  float* scratch;      // head-bias sums [16] | per time block: fresh bias grads of layer 0 / 5 [blocks][2][256] |
                       // tail: [0] amax of the incoming cotangents (as uint bits), [1] gscale, [2] 1/gscale
  float* dpe;          // [tiles*128, 96] dL/d(pe block) per row (general backward only)
  size_t scratch_floats;
  size_t total;
};

constexpr uint32_t kDoutBytes = kPanelBytes;

// A launch covers `blocks` time blocks of `rows` points each (blocks = 1: the plain [rows] call).  Every block
// starts on a tile boundary, so a tile never mixes two time values.
struct MlpShape {
  int64_t n;            // points per block
  int blocks;           // time values
  int tpb;              // tiles per block
  int tiles;            // blocks * tpb
};
inline MlpShape mlp_shape(int64_t n, int blocks) {
  MlpShape s;
  s.n = n; s.blocks = blocks < 1 ? 1 : blocks;
  s.tpb = (int)((n + kRows - 1) / kRows);
  if (s.tpb < 1) s.tpb = 1;
  s.tiles = s.tpb * s.blocks;
  return s;
}

inline MlpWs mlp_ws_view(void* base, int64_t rows, bool save, int blocks = 1) {
  MlpWs w;
  char* p = (char*)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { char* r = p + off; off += align_up(bytes, 1024); return r; };
  const MlpShape sh = mlp_shape(rows, blocks);
  size_t tiles = (size_t)sh.tiles;
  // time features: one row per point (plain call, rows may carry their own t) or one per time block
  size_t r = blocks > 1 ? (size_t)blocks : (size_t)(rows > 0 ? rows : 1);
  w.status = (int32_t*)take(1024);
  w.tfeat = (float*)take(r * 32 * 4);
  w.thidden = (float*)take(save ? r * 256 * 4 : 1024);
  if (save) {
    w.pe_img = (uint8_t*)take(tiles * kPeBytes);
    w.h_img = (uint8_t*)take(8 * tiles * (size_t)kHBytes);
    w.mask = (uint32_t*)take(8 * tiles * 8 * 128 * 4);
    w.heads_out = (float*)take(tiles * 128 * 16 * 4);
    w.dz_img = (uint8_t*)take(8 * tiles * (size_t)kHBytes);
    w.dout_img = (uint8_t*)take(tiles * (size_t)kDoutBytes);
    w.scratch_floats = 16 + (size_t)sh.blocks * 512 + 48;
    w.scratch = (float*)take(w.scratch_floats * 4);
    w.dpe = (float*)take(tiles * 128 * 96 * 4);
  } else {
    w.pe_img = w.h_img = w.dz_img = w.dout_img = nullptr;
    w.mask = nullptr; w.heads_out = nullptr; w.scratch = nullptr; w.dpe = nullptr;
    w.scratch_floats = 0;
  }
  w.total = off;
  return w;
}

// head index (row of the padded 16-wide head block) -> source parameter row
struct HeadMap {
  const float* w[16];
  const float* b[16];
  int r[16];
};
inline HeadMap head_map(const d3g_mlp_config& cfg, const d3g_mlp_params& p) {
  HeadMap m;
  for (int h = 0; h < 16; h++) {
    m.w[h] = nullptr; m.b[h] = nullptr; m.r[h] = 0;
    if (!cfg.is_6dof) {
      if (h < 3) { m.w[h] = p.warp_w; m.b[h] = p.warp_b; m.r[h] = h; }
      else if (h < 7) { if (cfg.heads & 2) { m.w[h] = p.rot_w; m.b[h] = p.rot_b; m.r[h] = h - 3; } }
      else if (h < 10) { if (cfg.heads & 4) { m.w[h] = p.scale_w; m.b[h] = p.scale_b; m.r[h] = h - 7; } }
    } else {
      if (h < 3) { m.w[h] = p.branch_w_w; m.b[h] = p.branch_w_b; m.r[h] = h; }
      else if (h < 6) { m.w[h] = p.branch_v_w; m.b[h] = p.branch_v_b; m.r[h] = h - 3; }
      else if (h < 10) { if (cfg.heads & 2) { m.w[h] = p.rot_w; m.b[h] = p.rot_b; m.r[h] = h - 6; } }
      else if (h < 13) { if (cfg.heads & 4) { m.w[h] = p.scale_w; m.b[h] = p.scale_b; m.r[h] = h - 10; } }
    }
  }
  return m;
}

int check_cfg(const d3g_mlp_config* cfg);
int num_sms();
int mlp_grid_limit();   // num_sms() minus the SMs reserved for a concurrent collective

}  // namespace d3g
