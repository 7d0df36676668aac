// Layout constants shared by the deformation-MLP kernels (forward, input-gradient chain, weight gradient).
//
// Network (utils/time_utils.py:56-201): PE(xyz) 63 | t-feature (timenet output 30, or PE(t) 21) -> "pe"
// block of pe_dim = 93 / 84 columns padded to 96; 8 layers of width 256, layer 5 takes [pe | h4];
// heads 256 -> up to 13 outputs padded to 16.
//
// One tile = 128 points.  All matrix operands are panelised SWIZZLE_128B images (umma.cuh).
#pragma once
#include "common.cuh"
#include "umma.cuh"

namespace d3g {

constexpr int kRows = 128;                 // points per tile
constexpr uint32_t kPanelBytes = 16384;    // 128 rows x 32 fp32
constexpr uint32_t kPeBytes = 3 * kPanelBytes;
constexpr uint32_t kHBytes = 8 * kPanelBytes;  // one 128x256 activation tile image
constexpr uint32_t kStageBytes = 32768;    // one weight stage: 256 rows x 32 k
constexpr int kFwdStages = 63;             // L0:3 L1-4:32 L5pe:3 L5h:8 L6-7:16 heads:1
constexpr int kBwdStages = 57;             // headsT:1 + 7 layers x 8
constexpr int kBiasFloats = 8 * 256 + 16;
constexpr size_t kFwdBytes = (size_t)62 * kStageBytes + 16384;
constexpr size_t kBiasOff = kFwdBytes;
constexpr size_t kBiasBytes = 8448;  // kBiasFloats*4 rounded up to 256
constexpr size_t kBwdOff = kBiasOff + kBiasBytes;
constexpr size_t kBwdBytes = (size_t)kBwdStages * kStageBytes;
constexpr size_t kPackedBytes = kBwdOff + kBwdBytes;

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
// Saved tiles are "MN-major wgrad images" (umma.cuh, layout (2)): exactly what the weight-gradient
// kernel feeds to the tensor core as operands with K = points, so backward never re-arranges them.
struct MlpWs {
  int32_t* status;     // [0] abort flag, [1] "rows carry different time values" (set on the device by forward)
  float* tfeat;        // [rows_t, 32]   time feature fed into pe columns 63..
  float* thidden;      // [rows_t, 256]  timenet hidden (post-ReLU), blender only
  uint8_t* pe_img;     // [tiles] x 48 KB            pe block (96 columns)
  uint8_t* h_img;      // [8][tiles] x 128 KB        post-ReLU activations h_0..h_7 (tf32-rounded)
  uint32_t* mask;      // [8][tiles][8 chunks][128 rows]  bit j = (h[row][chunk*32+j] > 0)
  float* heads_out;    // [tiles*128, 16] head pre-outputs (incl. bias)
  uint8_t* dz_img;     // [8][tiles] x 128 KB        dL/d(pre-activation) tiles, written by backward
  uint8_t* dout_img;   // [tiles] x 16 KB            head cotangents (32 columns, 16 used)
  float* scratch;      // head-bias sums [16] | per time block: fresh bias grads of layer 0 / 5 [blocks][2][256]
  float* dpe;          // [tiles*128, 96] dL/d(pe block) per row (general backward only)
  size_t scratch_floats;
  size_t total;
};

constexpr uint32_t kDoutBytes = 4 * 4096;

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

}  // namespace d3g
