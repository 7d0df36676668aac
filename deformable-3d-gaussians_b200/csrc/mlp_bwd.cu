// Deformation-MLP backward for sm_100a: input-gradient chain + weight gradients, both on tcgen05 (kind::f16 with
// fp32 accumulation, mlp_common.cuh).
//
// Reference: autograd through DeformNetwork(.Baseline).forward, utils/time_utils.py:102-127 / :176-201
// (8x256 trunk with the skip at layer 5, heads, optional exp_se3 of utils/rigid_utils.py:60-83).
//
// Range: gradients reach this code at any magnitude (a mean-reduced image loss puts them around 1e-7), which fp16
// cannot hold.  mlp_amax_kernel takes the largest |cotangent| of the call; every kernel derives the same power of two
// gscale from it, the chain runs on gscale * cotangent and the results are multiplied by 1/gscale (exact) on the way
// out.  Values more than ~2^-20 below the largest cotangent of the call are flushed — far below the parity bar, which
// is relative to each tensor's largest gradient.
//
// Kernel 1 (mlp_dgrad_kernel) — same persistent / warp-specialised shape as the forward: per 128-point
//   tile the head cotangents are gathered (and pushed back through the se(3) map for is_6dof) into a
//   shared-memory A operand, dH_7 = dOut * W_heads, then for l = 7..1: dZ_l = dH_l (.) relu'(h_l) is
//   formed IN PLACE in tensor memory (tcgen05.ld -> mask -> fp16 pack -> tcgen05.st) and consumed as the TMEM A
//   operand of dH_{l-1} = dZ_l * (W_l hi + W_l lo) — the weights are split (their rounding error would be the same
//   for every row and survive the sum over rows), dZ is not (its rounding is independent per row).  Each dZ_l tile is
//   also written to HBM as a [point][feature] fp16 image.
// Kernel 2 (mlp_wgrad_kernel) — dW_l = dZ_l^T * h_{l-1} is a GEMM whose K dimension is the points.
//   Both operands are read as MN-major images straight from what forward / kernel 1 saved; every CTA owns all 256
//   outputs of its tile range (two M=128 accumulators), the tile range is split-K over the CTAs, partial sums leave
//   TMEM through vector reductions (red.global.add.v4.f32) into the gradient tensors.  Bias gradients are
//   column sums of the dZ stages taken by the otherwise idle epilogue warps.
// Kernel 3 (mlp_finish_kernel) — head biases and the timenet (13->256->30) backward for a shared t.
#include <math.h>
#include <stdlib.h>

#include "mlp_common.cuh"

namespace d3g {

constexpr int kBwdNumStages = 5;
constexpr int kBwdThreads = 320;

struct BwdSmem {
  static constexpr uint32_t dout = 0;                       // 16 KB K-major panel (64 columns, 16 used)
  static constexpr uint32_t stages = kPanelBytes;
  static constexpr uint32_t store = stages + kBwdNumStages * kStageBytes;  // [4 lane quarters][2] x 4 KB bulk-store staging
  static constexpr uint32_t headb = store + 8 * 4096;                      // 16 floats
  static constexpr uint32_t bars = headb + 64;
  static constexpr uint32_t total = bars + 256;
};
static_assert(BwdSmem::bars % 8 == 0, "barrier alignment");

struct MlpBwdArgs {
  const uint8_t* wbwd;   // transposed weight stages
  int64_t rows;          // points per time block
  int tiles;             // all tiles (blocks * tpb)
  int tpb;               // tiles per time block
  int is_6dof, heads;
  const uint32_t* mask;
  const float* heads_out;
  const float* g_xyz;    // [rows,3] or [rows,4,4]
  const float* g_rot;    // [rows,4] or NULL
  const float* g_scale;  // [rows,3] or NULL
  uint8_t* dz_img;
  uint8_t* dout_img;
  float* headb_acc;      // [16] column sums of dOut
  uint32_t* amax_bits;
  const int* gain_e;
  int gtarget;
  int* abort_flag;
};

// Cotangent of the 16 head pre-outputs of one point.  Not 6dof: [g_xyz(3) | g_rot(4) | g_scale(3)].
// 6dof: [d w_raw(3) | d v_raw(3) | g_rot(4) | g_scale(3)], where (w_raw, v_raw) -> exp_se3 as in
// utils/time_utils.py:114-121 (w = w_raw/theta + 1e-5, v likewise, theta = |w_raw|).
__device__ __forceinline__ void head_cotangent(const MlpBwdArgs& a, int64_t grow, const float* ho, float (&o)[16]) {
#pragma unroll
  for (int j = 0; j < 16; j++) o[j] = 0.f;
  if (!a.is_6dof) {
    if (a.g_xyz) { o[0] = a.g_xyz[3 * grow]; o[1] = a.g_xyz[3 * grow + 1]; o[2] = a.g_xyz[3 * grow + 2]; }
    if (a.g_rot) { float4 r = *(const float4*)(a.g_rot + 4 * grow); o[3] = r.x; o[4] = r.y; o[5] = r.z; o[6] = r.w; }
    if (a.g_scale) { o[7] = a.g_scale[3 * grow]; o[8] = a.g_scale[3 * grow + 1]; o[9] = a.g_scale[3 * grow + 2]; }
    return;
  }
  if (a.g_rot) { float4 r = *(const float4*)(a.g_rot + 4 * grow); o[6] = r.x; o[7] = r.y; o[8] = r.z; o[9] = r.w; }
  if (a.g_scale) { o[10] = a.g_scale[3 * grow]; o[11] = a.g_scale[3 * grow + 1]; o[12] = a.g_scale[3 * grow + 2]; }
  if (!a.g_xyz) return;
  const float4* G4 = (const float4*)(a.g_xyz + 16 * grow);
  float4 g0 = G4[0], g1 = G4[1], g2 = G4[2];
  float GR[9] = {g0.x, g0.y, g0.z, g1.x, g1.y, g1.z, g2.x, g2.y, g2.z};
  float Gp[3] = {g0.w, g1.w, g2.w};
  float wr[3] = {ho[0], ho[1], ho[2]}, vr[3] = {ho[3], ho[4], ho[5]};
  float theta = sqrtf(wr[0] * wr[0] + wr[1] * wr[1] + wr[2] * wr[2]);
  float it = 1.0f / theta;
  float w[3] = {wr[0] * it + 1e-5f, wr[1] * it + 1e-5f, wr[2] * it + 1e-5f};
  float v[3] = {vr[0] * it + 1e-5f, vr[1] * it + 1e-5f, vr[2] * it + 1e-5f};
  float sn, cs;
  sincosf(theta, &sn, &cs);
  float W[9] = {0.f, -w[2], w[1], w[2], 0.f, -w[0], -w[1], w[0], 0.f};
  float W2[9];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) W2[3 * i + j] = W[3 * i] * W[j] + W[3 * i + 1] * W[3 + j] + W[3 * i + 2] * W[6 + j];
  const float c1 = 1.0f - cs, c2 = theta - sn;
  // p = P v, P = theta I + c1 W + c2 W2;  R = I + sn W + c1 W2
  float dP[9], dv[3] = {0.f, 0.f, 0.f};
  float dtheta = 0.f;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      float id = i == j ? 1.f : 0.f;
      float Pij = theta * id + c1 * W[3 * i + j] + c2 * W2[3 * i + j];
      dv[j] += Pij * Gp[i];
      dP[3 * i + j] = Gp[i] * v[j];
      // dR/dtheta = cs W + sn W2 ; dP/dtheta = I + sn W + c1 W2
      dtheta += GR[3 * i + j] * (cs * W[3 * i + j] + sn * W2[3 * i + j]) +
                dP[3 * i + j] * (id + sn * W[3 * i + j] + c1 * W2[3 * i + j]);
    }
  // cotangent of W2 and of W
  float dW2[9], dW[9];
#pragma unroll
  for (int i = 0; i < 9; i++) { dW2[i] = c1 * GR[i] + c2 * dP[i]; dW[i] = sn * GR[i] + c1 * dP[i]; }
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 3; k++) acc += dW2[3 * i + k] * W[3 * j + k] + W[3 * k + i] * dW2[3 * k + j];
      dW[3 * i + j] += acc;  // G W^T + W^T G
    }
  float dw[3] = {dW[7] - dW[5], dW[2] - dW[6], dW[3] - dW[1]};
  float dot = 0.f;
#pragma unroll
  for (int d = 0; d < 3; d++) dot += dw[d] * wr[d] + dv[d] * vr[d];
  dtheta -= dot * it * it;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    o[d] = dw[d] * it + dtheta * wr[d] * it;
    o[3 + d] = dv[d] * it;
  }
}

// largest |head cotangent| of the call (after the se(3) pull-back for 6-DoF): one thread per row
__global__ void __launch_bounds__(256) mlp_amax_kernel(MlpBwdArgs a) {
  float m = 0.f;
  const int64_t total = (int64_t)(a.tiles / a.tpb) * a.rows;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int blk = (int)(i / a.rows);
    const int64_t prow = i - (int64_t)blk * a.rows;
    float ho[16], o[16];
    if (a.is_6dof) {
      const float4* h4 = (const float4*)(a.heads_out + ((size_t)blk * a.tpb * kRows + prow) * 16);
      float4 h0 = h4[0], h1 = h4[1];
      ho[0] = h0.x; ho[1] = h0.y; ho[2] = h0.z; ho[3] = h0.w; ho[4] = h1.x; ho[5] = h1.y;
    }
    head_cotangent(a, i, ho, o);
#pragma unroll
    for (int j = 0; j < 16; j++) m = fmaxf(m, fabsf(o[j]));
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(a.amax_bits, __float_as_uint(m));
}

__device__ __forceinline__ void stg128b(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(kBwdThreads, 1) mlp_dgrad_kernel(MlpBwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sDout = smem + BwdSmem::dout;
  uint8_t* sStage = smem + BwdSmem::stages;
  float* sHeadB = (float*)(smem + BwdSmem::headb);
  uint8_t* sStore = smem + BwdSmem::store;
  uint64_t* bars = (uint64_t*)(smem + BwdSmem::bars);
  uint64_t* full = bars;                      // [kBwdNumStages]
  uint64_t* empty = bars + kBwdNumStages;     // [kBwdNumStages]
  uint64_t* acc_full = bars + 2 * kBwdNumStages;
  uint64_t* dout_ready = acc_full + 1;
  // go[c]: chunk c of the next chain step may be multiplied = handed over by the four epilogue warps that produce it AND
  // its weight stage has landed (same scheme as the forward kernel: one barrier poll per chunk on the issue thread)
  uint64_t* go = dout_ready + 1;              // [8]
  uint32_t* tmem_slot = (uint32_t*)(go + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  volatile int* abort_flag = a.abort_flag;

  if (tid == 0) {
    for (int i = 0; i < kBwdNumStages; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_full, 1);
    mbar_init(dout_ready, 8);
    for (int i = 0; i < 8; i++) mbar_init(&go[i], 5);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(tmem_slot, 512);
  if (tid < 16) sHeadB[tid] = 0.f;
  // the dOut panel is only ever written in its first 16 logical columns; clear it once
  for (int i = tid; i < (int)(kPanelBytes / 16); i += kBwdThreads) ((uint4*)sDout)[i] = make_uint4(0, 0, 0, 0);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ================= weight producer (transposed stages) =================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      bool ok = true;
      for (int tile = blockIdx.x; ok && tile < a.tiles; tile += gridDim.x) {
        for (int st = 0; st < kBwdStages; st++) {
          if (!mbar_wait(&empty[s], ph ^ 1, abort_flag)) { ok = false; break; }
          uint64_t* bar = st == 0 ? &full[s] : &go[(st - 1) & 7];   // stage 0 = heads^T, then 7 x 8 chunk stages
          mbar_arrive_expect_tx(bar, kStageBytes);
          bulk_g2s(sStage + (size_t)s * kStageBytes, a.wbwd + (size_t)st * kStageBytes, kStageBytes, bar);
          if (++s == kBwdNumStages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 9) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc256 = make_idesc_f16(128, 256, 0, 0);
      int s = 0;
      uint32_t ph = 0, d_ph = 0, a_ph = 0, full_bits = 0;   // (full[s] only completes for the heads^T stage: per-slot parity)
      bool ok = true;
      const uint64_t bdesc0 = make_smem_desc(smem_u32(sStage), 16, 1024);
      auto release = [&]() { umma_commit(&empty[s]); if (++s == kBwdNumStages) { s = 0; ph ^= 1; } };
      for (int tile = blockIdx.x; ok && tile < a.tiles; tile += gridDim.x) {
        // step 0: dH_7 = dOut[128,16] * Wheads[16,256]  (A from shared memory, K = 16; W = hi + lo)
        if (!mbar_wait(dout_ready, d_ph, abort_flag)) { ok = false; break; }
        d_ph ^= 1;
        if (!mbar_wait(&full[s], (full_bits >> s) & 1u, abort_flag)) { ok = false; break; }
        full_bits ^= 1u << s;
        tc_fence_after();
        {
          const uint64_t sb = desc_adv(bdesc0, (uint32_t)s * kStageBytes);
          const uint64_t ad = make_smem_desc(smem_u32(sDout), 16, 1024);
          umma_f16_ss(tmem_base + 0, ad, sb, idesc256, 0u);
          umma_f16_ss(tmem_base + 0, ad, desc_adv(sb, 64), idesc256, 1u);
          release();
          umma_commit(acc_full);
        }
        // steps 1..7: dH_{7-j} = dZ_{8-j} * W_{8-j}; A = TMEM half (j-1)&1 (packed fp16), D = half j&1
        for (int j = 1; ok && j < 8; j++) {
          const uint32_t d_col = (j & 1) * 256, a_col = ((j - 1) & 1) * 256;
          for (int kc = 0; ok && kc < 8; kc++) {
            if (!mbar_wait(&go[kc], a_ph, abort_flag)) { ok = false; break; }
            tc_fence_after();
            const uint64_t sb = desc_adv(bdesc0, (uint32_t)s * kStageBytes);
            const uint32_t ta = tmem_base + a_col + kc * 32;
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
              umma_f16_ts(tmem_base + d_col, ta + ks * 8, desc_adv(sb, ks * 32), idesc256, (kc | ks) != 0);
              umma_f16_ts(tmem_base + d_col, ta + ks * 8, desc_adv(sb, 64 + ks * 32), idesc256, 1u);
            }
            release();
          }
          a_ph ^= 1;
          if (ok) umma_commit(acc_full);
        }
      }
    }
  } else {
    // ================= cotangent gather + epilogue warps (256 threads) =================
    const int q = warp & 3, chalf = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t acc_ph = 0;
    bool ok = true;
    const GradScales G = grad_scales(*a.amax_bits, a.gtarget, a.gain_e);
    const float gs = G.gs;
    for (int tile = blockIdx.x; ok && tile < a.tiles; tile += gridDim.x) {
      const int blk = tile / a.tpb;
      const int64_t prow = (int64_t)(tile - blk * a.tpb) * kRows + row;
      const bool live = prow < a.rows;
      const int64_t grow = (int64_t)blk * a.rows + prow;   // row of the [blocks, rows, .] cotangents
      // ---- head cotangents (x gscale, fp16) -> shared A panel + HBM image; column sums for the head biases ----
      if (chalf == 0) {
        float o[16];
        if (live) {
          float ho[16];
          if (a.is_6dof) {
            const float4* h4 = (const float4*)(a.heads_out + ((size_t)tile * kRows + row) * 16);
            float4 h0 = h4[0], h1 = h4[1];
            ho[0] = h0.x; ho[1] = h0.y; ho[2] = h0.z; ho[3] = h0.w; ho[4] = h1.x; ho[5] = h1.y;
          }
          head_cotangent(a, grow, ho, o);
        } else {
#pragma unroll
          for (int j = 0; j < 16; j++) o[j] = 0.f;
        }
        uint32_t u[8];
#pragma unroll
        for (int j = 0; j < 8; j++) u[j] = pack2(o[2 * j] * gs, o[2 * j + 1] * gs);
        uint8_t* dst = a.dout_img + (size_t)tile * kDoutBytes + row * 128;
#pragma unroll
        for (int g = 0; g < 2; g++) {
          sts128(smem_u32(sDout) + row * 128 + ((g ^ (row & 7)) << 4), u[4 * g], u[4 * g + 1], u[4 * g + 2], u[4 * g + 3]);
          stg128b(dst + ((g ^ (row & 7)) << 4), u[4 * g], u[4 * g + 1], u[4 * g + 2], u[4 * g + 3]);
        }
#pragma unroll
        for (int j = 0; j < 16; j++) {
          float sum = o[j];
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
          if (lane == 0) atomicAdd(&sHeadB[j], sum);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(dout_ready);

      // ---- 8 epilogues: dZ_l = dH_l (.) relu'(h_l), l = 7..0 ----
      for (int j = 0; ok && j < 8; j++) {
        const int l = 7 - j;
        uint32_t mbits[4];
#pragma unroll
        for (int i = 0; i < 4; i++)
          mbits[i] = a.mask[(((size_t)l * a.tiles + tile) * 8 + (chalf + 2 * i)) * 128 + row];
        {   // one lane polls, the warp follows
          int got = 1;
          if (lane == 0) got = mbar_wait(acc_full, acc_ph, abort_flag) ? 1 : 0;
          got = __shfl_sync(0xffffffffu, got, 0);
          if (!got) { ok = false; break; }
        }
        acc_ph ^= 1;
        tc_fence_after();
        const uint32_t half = (j & 1) * 256;
        const float mul = G.step_mul[j];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int c = chalf + 2 * i;
          uint32_t v[32];
          tmem_ld32(lane_addr + half + c * 32, v);
          tmem_wait_ld();
          const uint32_t bits = mbits[i];
          uint32_t o[16];
#pragma unroll
          for (int k = 0; k < 32; k += 2) {
            const float x0 = (bits & (1u << k)) ? __uint_as_float(v[k]) * mul : 0.f;
            const float x1 = (bits & (2u << k)) ? __uint_as_float(v[k + 1]) * mul : 0.f;
            o[k >> 1] = pack2(x0, x1);
          }
          if (j < 7) {
            tmem_st16(lane_addr + half + c * 32, o);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&go[c]);
          }
          // dZ -> [point][feature] image: the two warps of a lane quarter assemble the 4 KB block (32 rows of panel
          // c/2) in shared memory and the bulk-copy engine stores it (see the forward kernel)
          {
            const int pr = c >> 1;
            uint8_t* stg = sStore + ((q * 2 + (pr & 1)) << 12);
            if (chalf == 0 && lane == 0) bulk_wait_read<1>();
            asm volatile("bar.sync %0, %1;" ::"r"(1 + q), "r"(64) : "memory");
            const uint32_t line = smem_u32(stg) + lane * 128;
#pragma unroll
            for (int g = 0; g < 4; g++)
              sts128(line + (((4 * (c & 1) + g) ^ (lane & 7)) << 4), o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
            fence_proxy_async();
            asm volatile("bar.sync %0, %1;" ::"r"(1 + q), "r"(64) : "memory");
            if (chalf == 0 && lane == 0) {
              bulk_s2g(a.dz_img + ((size_t)l * a.tiles + tile) * kHBytes + (size_t)pr * kPanelBytes + (size_t)q * 4096, stg, 4096);
              bulk_commit();
            }
          }
        }
      }
      tc_fence_before();
    }
    if (chalf == 0 && lane == 0) bulk_wait_all<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (tid < 16) atomicAdd(&a.headb_acc[tid], sHeadB[tid]);
  if (warp == 9) tmem_dealloc(tmem_base, 512);
}

// ================================================================================================
// Kernel 2: weight gradients.  One work item = one (A image set, B image set) GEMM over all tiles.
// A stage holds 64 points: A = dZ slice [4 feature panels][64 rows][128 B] (32 KB), B = activation slice
// [N/64 panels][64 rows][128 B]; both are MN-major SWIZZLE_128B operands (LBO = 8 KB panel stride, SBO = 1 KB).
// ================================================================================================
constexpr int kWgStages = 3;
constexpr int kWgPoints = 64;                         // points per stage
constexpr uint32_t kWgPanel = kWgPoints * 128;        // 8 KB: one 64-feature panel of a stage
constexpr uint32_t kWgABytes = 4 * kWgPanel;          // 64 points x 256 features (both halves of the outputs)
constexpr uint32_t kWgBBytes = 4 * kWgPanel;          // 64 points x up to 256 features
constexpr uint32_t kWgStageBytes = kWgABytes + kWgBBytes;
constexpr int kWgThreads = 192;                       // warp 0 producer, 1 MMA, 2..5 epilogue
constexpr int kWgItems = 10;

struct WgItem {
  const uint8_t* a_img;   // [tiles] x 64 KB images of 256 features (dz_l, or h_7 for the heads)
  const uint8_t* b_img;   // [tiles] x b_tile_bytes
  uint32_t b_tile_bytes;  // 65536 (h), 32768 (pe), 16384 (dout)
  int b_panels;           // 64-feature panels of B to load: 4, 2 or 1
  int N;                  // 256, 96 or 16
  int bias;               // column sums of the A stages -> bias gradient
  int kind;               // 0: trunk layer (dW[o][col0+i]), 1: heads (transposed scatter)
  int layer;              // trunk: whose dZ image A is (its scale 2^E_layer is undone on the way out); -1: heads
  int cost;               // relative cost per tile: KB moved (the kernel is HBM-bound)
  float* dW;              // trunk: gradient of linear.l.weight
  int ld, col0, ncols;
  float* db;              // trunk: gradient of linear.l.bias (or NULL)
  float* db_fresh;        // scratch copy of this call's bias gradient (layers 0 and 5) or NULL
};

struct MlpWgradArgs {
  int dbg;                // bit0: skip the bulk loads, bit1: skip the MMAs (timing experiments only)
  WgItem items[kWgItems];
  int n_items;
  int tiles;
  int tpb;                // tiles per time block: db_fresh is kept per block ([blocks][2][256])
  float* head_dw[16];     // heads: gradient row pointers (NULL where the head is not evaluated)
  const uint32_t* amax_bits;
  const int* gain_e;
  int gtarget;
  int* abort_flag;
};

// tile range of item `it` that pair `pair` (of `pairs`) owns; ranges of consecutive pairs tile the item
__device__ __forceinline__ void wg_segment(const MlpWgradArgs& a, int pair, int pairs, int it, int64_t base,
                                           int64_t total, int& t0, int& t1) {
  const int64_t lo = total * pair / pairs, hi = total * (pair + 1) / pairs;
  const int64_t c = a.items[it].cost;
  const int64_t ib = base, ie = base + c * a.tiles;
  const int64_t s = lo > ib ? lo : ib, e = hi < ie ? hi : ie;
  if (s >= e) { t0 = t1 = 0; return; }
  t0 = (int)((s - ib + c - 1) / c);
  t1 = (int)((e - ib + c - 1) / c);
}

__device__ __forceinline__ void red_add_v4f(float* addr, float x, float y, float z, float w) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

__global__ void __launch_bounds__(kWgThreads, 1) mlp_wgrad_kernel(MlpWgradArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sStage = smem;                                   // kWgStages x 64 KB
  uint64_t* bars = (uint64_t*)(smem + kWgStages * kWgStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + kWgStages;
  uint64_t* acc_full = bars + 2 * kWgStages;
  uint64_t* acc_empty = acc_full + 1;
  uint32_t* tmem_slot = (uint32_t*)(acc_empty + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  volatile int* abort_flag = a.abort_flag;
  // every CTA owns all 256 outputs of its tile range (two M=128 accumulators = all 512 TMEM columns): the B stage
  // is then read once per point instead of once per CTA of a pair
  const int pair = blockIdx.x, pairs = gridDim.x;

  if (tid == 0) {
    for (int i = 0; i < kWgStages; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 5); }  // MMA commit + 4 warps
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 4);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  int64_t total = 0;
  for (int it = 0; it < a.n_items; it++) total += (int64_t)a.items[it].cost * a.tiles;

  if (warp == 0) {
    // ================= producer: 64 points of A (all 256 features) and B per stage ==========
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      bool ok = true;
      int64_t base = 0;
      for (int it = 0; ok && it < a.n_items; it++) {
        const WgItem& w = a.items[it];
        int t0, t1;
        wg_segment(a, pair, pairs, it, base, total, t0, t1);
        base += (int64_t)w.cost * a.tiles;
        for (int tile = t0; ok && tile < t1; tile++) {
          const uint8_t* pa = w.a_img + (size_t)tile * kHBytes;
          const uint8_t* pb = w.b_img + (size_t)tile * w.b_tile_bytes;
          for (int ch = 0; ch < kRows / kWgPoints; ch++) {
            if (!mbar_wait(&empty[s], ph ^ 1, abort_flag)) { ok = false; break; }
            uint8_t* dst = sStage + (size_t)s * kWgStageBytes;
            if (a.dbg & 1) { mbar_arrive(&full[s]); }
            else {
              mbar_arrive_expect_tx(&full[s], kWgABytes + (uint32_t)w.b_panels * kWgPanel);
#pragma unroll
              for (int pn = 0; pn < 4; pn++)
                bulk_g2s(dst + pn * kWgPanel, pa + (size_t)pn * kPanelBytes + (size_t)ch * kWgPanel, kWgPanel, &full[s]);
              for (int pn = 0; pn < w.b_panels; pn++)
                bulk_g2s(dst + kWgABytes + pn * kWgPanel, pb + (size_t)pn * kPanelBytes + (size_t)ch * kWgPanel, kWgPanel,
                         &full[s]);
            }
            if (++s == kWgStages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0, e_ph = 0;
      bool ok = true;
      int64_t base = 0;
      for (int it = 0; ok && it < a.n_items; it++) {
        const WgItem& w = a.items[it];
        int t0, t1;
        wg_segment(a, pair, pairs, it, base, total, t0, t1);
        base += (int64_t)w.cost * a.tiles;
        if (t0 >= t1) continue;
        const uint32_t idesc = make_idesc_f16(128, (uint32_t)w.N, 1, 1);
        // the previous segment's accumulator must have been drained
        if (!mbar_wait(acc_empty, e_ph ^ 1, abort_flag)) { ok = false; break; }
        tc_fence_after();
        uint32_t first = 1;
        for (int tile = t0; ok && tile < t1; tile++) {
          for (int ch = 0; ch < kRows / kWgPoints; ch++) {
            if (!mbar_wait(&full[s], ph, abort_flag)) { ok = false; break; }
            tc_fence_after();
            const uint32_t sa = smem_u32(sStage + (size_t)s * kWgStageBytes), sb = sa + kWgABytes;
#pragma unroll
            for (int k16 = 0; k16 < kWgPoints / 16; k16++) {
              if (a.dbg & 2) break;
              const uint64_t bd = make_smem_desc_t(sb + k16 * 2048, kWgPanel, 1024, 2u);
              umma_f16_ss(tmem_base, make_smem_desc_t(sa + k16 * 2048, kWgPanel, 1024, 2u), bd, idesc, first ? 0u : 1u);
              umma_f16_ss(tmem_base + 256, make_smem_desc_t(sa + 2 * kWgPanel + k16 * 2048, kWgPanel, 1024, 2u), bd, idesc,
                          first ? 0u : 1u);
              first = 0;
            }
            if (a.dbg & 8) mbar_arrive(&empty[s]); else umma_commit(&empty[s]);
            if (++s == kWgStages) { s = 0; ph ^= 1; }
          }
        }
        if (ok) { umma_commit(acc_full); e_ph ^= 1; }
      }
    }
  } else {
    // ================= epilogue: TMEM -> reductions into the gradient tensors =================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const uint32_t lane_addr0 = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t f_ph = 0;
    bool ok = true;
    int64_t base = 0;
    int s = 0;
    uint32_t ph = 0;
    const GradScales G = grad_scales(*a.amax_bits, a.gtarget, a.gain_e);
    // bias gradients: thread j sums features 2j, 2j+1 over the 64 points of every A stage
    const int f = 2 * ((warp - 2) * 32 + lane);
    const uint32_t f_off = (uint32_t)(f >> 6) * kWgPanel + (uint32_t)(f & 7) * 2u, f_chunk = (uint32_t)(f & 63) >> 3;
    for (int it = 0; ok && it < a.n_items; it++) {
      const WgItem& w = a.items[it];
      int t0, t1;
      wg_segment(a, pair, pairs, it, base, total, t0, t1);
      base += (int64_t)w.cost * a.tiles;
      if (t0 >= t1) continue;
      const float gs_inv = w.layer >= 0 ? G.dz_inv[w.layer] : G.gs_inv;
      float bsum = 0.f, bsum1 = 0.f;
      int cur_blk = t0 / a.tpb;
      auto flush_bias = [&]() {
        if (w.db) { atomicAdd(w.db + f, bsum * gs_inv); atomicAdd(w.db + f + 1, bsum1 * gs_inv); }
        if (w.db_fresh) {
          atomicAdd(w.db_fresh + (size_t)cur_blk * 512 + f, bsum * gs_inv);
          atomicAdd(w.db_fresh + (size_t)cur_blk * 512 + f + 1, bsum1 * gs_inv);
        }
        bsum = bsum1 = 0.f;
      };
      for (int tile = t0; ok && tile < t1; tile++) {
        if (w.bias && tile / a.tpb != cur_blk) {   // next time block: flush this block's bias sums
          flush_bias();
          cur_blk = tile / a.tpb;
        }
        for (int ch = 0; ch < kRows / kWgPoints; ch++) {
          if (!mbar_wait(&full[s], ph, abort_flag)) { ok = false; break; }
          if (w.bias) {
            const uint32_t sa = smem_u32(sStage + (size_t)s * kWgStageBytes) + f_off;
            float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
#pragma unroll
            for (int p = 0; p < kWgPoints; p += 2) {
              uint32_t u0, u1;
              asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u0) : "r"(sa + p * 128 + ((f_chunk ^ (p & 7)) << 4)));
              asm volatile("ld.shared.b32 %0, [%1];" : "=r"(u1) : "r"(sa + (p + 1) * 128 + ((f_chunk ^ ((p + 1) & 7)) << 4)));
              const float2 x0 = __half22float2(*(const __half2*)&u0), x1 = __half22float2(*(const __half2*)&u1);
              acc0 += x0.x; acc1 += x0.y; acc2 += x1.x; acc3 += x1.y;
            }
            bsum += acc0 + acc2;
            bsum1 += acc1 + acc3;
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty[s]);
          if (++s == kWgStages) { s = 0; ph ^= 1; }
        }
      }
      if (!ok) break;
      if (!mbar_wait(acc_full, f_ph, abort_flag)) { ok = false; break; }
      f_ph ^= 1;
      tc_fence_after();
      for (int hh = 0; hh < 2; hh++) {
        const int o = hh * 128 + q * 32 + lane;
        const uint32_t lane_addr = lane_addr0 + hh * 256;
        if (w.kind == 0) {
          for (int c0 = 0; c0 < w.N; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(lane_addr + c0, v);
            tmem_wait_ld();
            if (a.dbg & 4) continue;
            float* dst = w.dW + (size_t)o * w.ld + w.col0 + c0;
#pragma unroll
            for (int j = 0; j < 8; j++) {
              float* p4 = dst + 4 * j;
              const float x0 = __uint_as_float(v[4 * j]) * gs_inv, x1 = __uint_as_float(v[4 * j + 1]) * gs_inv,
                          x2 = __uint_as_float(v[4 * j + 2]) * gs_inv, x3 = __uint_as_float(v[4 * j + 3]) * gs_inv;
              if (c0 + 4 * j + 3 < w.ncols && (((uintptr_t)p4) & 15) == 0) {
                red_add_v4f(p4, x0, x1, x2, x3);
              } else {
                if (c0 + 4 * j < w.ncols) atomicAdd(p4, x0);
                if (c0 + 4 * j + 1 < w.ncols) atomicAdd(p4 + 1, x1);
                if (c0 + 4 * j + 2 < w.ncols) atomicAdd(p4 + 2, x2);
                if (c0 + 4 * j + 3 < w.ncols) atomicAdd(p4 + 3, x3);
              }
            }
          }
        } else {
          // heads: D[in-feature o][head j] -> d(head weight row j)[o]
          uint32_t v[16];
          tmem_ld16(lane_addr, v);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; j++)
            if (a.head_dw[j]) atomicAdd(a.head_dw[j] + o, __uint_as_float(v[j]) * gs_inv);
        }
      }
      if (w.bias) flush_bias();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ================================================================================================
// Kernel 3: head-bias gradients, and the timenet backward of every time block whose rows share one time value.
//   t-feature columns 63..92 of the pe block feed layer 0 and the skip layer, so with a shared t
//   sum_rows d tfeat[j] = sum_o db_0[o] W_0[o][63+j] + db_5[o] W_5[o][63+j]   (db = this block's bias gradients),
//   then utils/time_utils.py:74-76 backwards: tfeat = W2 relu(W0t e + b0t) + b2.  Everything behind tfeat is
//   constant over the block, so the same sums give sum_rows dL/dt (dL_dt_blocks, nullable).
//   One CTA per time block; parameter gradients are accumulated with atomics.
// ================================================================================================
struct MlpFinishArgs {
  const float* headb_acc;   // [16]
  float* head_db[16];       // gradient element pointers (NULL where absent)
  int is_blender;
  int per_row;              // plain call with per-row time values: when they differ, mlp_timenet_rows_kernel
  const int* nonuniform;    // owns the timenet gradients and this kernel only does the head biases
  const float* db_fresh;    // [blocks][2][256]
  const float* w0;          // linear.0.weight [256, pe_dim]
  const float* w5;          // linear.5.weight [256, pe_dim+256]
  int pe_dim;
  const float* t;           // time value of block b at t[b * t_step]
  int t_step;
  const float* thidden;     // [blocks][256] post-ReLU timenet hidden
  const float* tn0_w;       // timenet.0.weight [256,13]
  const float* tn2_w;       // timenet.2.weight [30,256]
  float* g_tn0_w; float* g_tn0_b; float* g_tn2_w; float* g_tn2_b;
  float* dL_dt_blocks;      // [blocks] or NULL
};

__global__ void __launch_bounds__(256) mlp_finish_kernel(MlpFinishArgs a) {
  __shared__ float dtf[32];
  __shared__ float e[32];
  __shared__ float de[16];
  const int tid = threadIdx.x, blk = blockIdx.x;
  if (blk == 0 && tid < 16 && a.head_db[tid]) atomicAdd(a.head_db[tid], a.headb_acc[tid]);
  if (a.per_row && *a.nonuniform) return;   // the per-row kernels handle the time path
  if (!a.is_blender && !a.dL_dt_blocks) return;
  const int L = a.is_blender ? 6 : 10, t_dim = a.is_blender ? 30 : 21;
  if (tid < 32) {
    const float tv = a.t[(size_t)blk * a.t_step];
    float v = 0.f;
    if (tid == 0) v = tv;
    else if (tid < 1 + 2 * L) {
      const int f = (tid - 1) >> 1;
      const float arg = tv * (float)(1 << f);
      v = ((tid - 1) & 1) ? cosf(arg) : sinf(arg);
    }
    e[tid] = v;
  }
  if (tid < 16) de[tid] = 0.f;
  {
    const int warp = tid >> 5, lane = tid & 31;
    const float* dbf = a.db_fresh + (size_t)blk * 512;
    for (int j = warp; j < t_dim; j += 8) {
      float acc = 0.f;
#pragma unroll
      for (int o = lane; o < 256; o += 32)
        acc += dbf[o] * a.w0[(size_t)o * a.pe_dim + 63 + j] + dbf[256 + o] * a.w5[(size_t)o * (a.pe_dim + 256) + 63 + j];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
      if (lane == 0) dtf[j] = acc;
    }
  }
  __syncthreads();
  if (!a.is_blender) {
    // tfeat = PE_10(t) itself: d/dt of [t, sin(2^f t), cos(2^f t)]
    if (tid == 0) {
      float dt = dtf[0];
      for (int f = 0; f < L; f++) dt += (dtf[1 + 2 * f] * e[2 + 2 * f] - dtf[2 + 2 * f] * e[1 + 2 * f]) * (float)(1 << f);
      a.dL_dt_blocks[blk] = dt;
    }
    return;
  }
  // thread k owns hidden unit k
  const int k = tid;
  const float hk = a.thidden[(size_t)blk * 256 + k];
  float dh = 0.f;
#pragma unroll
  for (int i = 0; i < 30; i++) {
    atomicAdd(&a.g_tn2_w[i * 256 + k], dtf[i] * hk);
    dh += a.tn2_w[i * 256 + k] * dtf[i];
  }
  if (k < 30) atomicAdd(&a.g_tn2_b[k], dtf[k]);
  if (!(hk > 0.f)) dh = 0.f;
  atomicAdd(&a.g_tn0_b[k], dh);
#pragma unroll
  for (int c = 0; c < 13; c++) atomicAdd(&a.g_tn0_w[k * 13 + c], dh * e[c]);
  if (a.dL_dt_blocks) {
#pragma unroll
    for (int c = 0; c < 13; c++) {
      float v = dh * a.tn0_w[k * 13 + c];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
      if ((tid & 31) == 0) atomicAdd(&de[c], v);
    }
    __syncthreads();
    if (tid == 0) {
      float dt = de[0];
      for (int f = 0; f < 6; f++) dt += (de[1 + 2 * f] * e[2 + 2 * f] - de[2 + 2 * f] * e[1 + 2 * f]) * (float)(1 << f);
      a.dL_dt_blocks[blk] = dt;
    }
  }
}

// ================================================================================================
// General backward of the input block (CUDA cores; only launched when a caller wants dL/dxyz or dL/dt, or when the
// rows of a blender call carry different time values):
//   dPE[p][c] = sum_o dZ_0[p][o] W_0[o][c] + dZ_5[p][o] W_5[o][c],  c < pe_dim        (mlp_dpe_kernel)
//   dL/dxyz through [x, sin(2^f x), cos(2^f x)] (utils/time_utils.py:46-54)            (mlp_pe_bwd_kernel)
//   per-row timenet backward 30 <- 256 <- 13 and dL/dt                                 (mlp_timenet_rows_kernel)
// ================================================================================================
struct MlpDpeArgs {
  const uint8_t* dz0;   // gscale * dZ images (fp16) of layer 0 and layer 5, [tiles] x 64 KB
  const uint8_t* dz5;
  const float* w0;      // [256, pe_dim]
  const float* w5;      // [256, pe_dim + 256]
  int pe_dim;
  int tiles;
  float* dpe;           // [tiles*128, 96]
  int skip_if_uniform;  // nothing downstream needs dPE when the time values turned out equal
  const int* nonuniform;
  const uint32_t* amax_bits;
  const int* gain_e;
  int gtarget;
};

__global__ void __launch_bounds__(256) mlp_dpe_kernel(MlpDpeArgs a) {
  __shared__ __align__(16) uint8_t sA[kPanelBytes];   // 128 points x 64 features, one image panel as stored
  __shared__ __align__(16) float sB[64 * 96];
  if (a.skip_if_uniform && !*a.nonuniform) return;
  const GradScales G = grad_scales(*a.amax_bits, a.gtarget, a.gain_e);
  const int tid = threadIdx.x, rg = tid >> 4, cgp = tid & 15;
  for (int tile = blockIdx.x; tile < a.tiles; tile += gridDim.x) {
    float acc[8][6], tot[8][6];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int j = 0; j < 6; j++) tot[i][j] = 0.f;
    for (int src = 0; src < 2; src++) {
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 6; j++) acc[i][j] = 0.f;
      const uint8_t* img = (src ? a.dz5 : a.dz0) + (size_t)tile * kHBytes;
      const float* W = src ? a.w5 : a.w0;
      const int ld = src ? a.pe_dim + 256 : a.pe_dim;
      for (int kc = 0; kc < 4; kc++) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; i++)
          ((float4*)sA)[tid + i * 256] = *(const float4*)(img + (size_t)kc * kPanelBytes + (size_t)(tid + i * 256) * 16);
        for (int idx = tid; idx < 64 * 96; idx += 256) {
          const int k = idx / 96, c = idx - k * 96;
          sB[idx] = c < a.pe_dim ? W[(size_t)(kc * 64 + k) * ld + c] : 0.f;
        }
        __syncthreads();
#pragma unroll 2
        for (int k8 = 0; k8 < 8; k8++) {
          float av[8][8];
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const int p = rg * 8 + i;
            const uint4 u = *(const uint4*)(sA + p * 128 + ((k8 ^ (p & 7)) << 4));
            const float2 f0 = __half22float2(*(const __half2*)&u.x), f1 = __half22float2(*(const __half2*)&u.y),
                         f2 = __half22float2(*(const __half2*)&u.z), f3 = __half22float2(*(const __half2*)&u.w);
            av[i][0] = f0.x; av[i][1] = f0.y; av[i][2] = f1.x; av[i][3] = f1.y;
            av[i][4] = f2.x; av[i][5] = f2.y; av[i][6] = f3.x; av[i][7] = f3.y;
          }
#pragma unroll
          for (int kk = 0; kk < 8; kk++) {
            const float* br = sB + (k8 * 8 + kk) * 96 + cgp * 6;
            const float2 b0 = *(const float2*)br, b1 = *(const float2*)(br + 2), b2 = *(const float2*)(br + 4);
            const float bv[6] = {b0.x, b0.y, b1.x, b1.y, b2.x, b2.y};
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
              for (int j = 0; j < 6; j++) acc[i][j] = fmaf(av[i][kk], bv[j], acc[i][j]);
          }
        }
      }
      const float inv = G.dz_inv[src ? 5 : 0];
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 6; j++) tot[i][j] = fmaf(acc[i][j], inv, tot[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 8; i++) {
      float* dst = a.dpe + ((size_t)tile * kRows + rg * 8 + i) * 96 + cgp * 6;
#pragma unroll
      for (int j = 0; j < 3; j++) *(float2*)(dst + 2 * j) = make_float2(tot[i][2 * j], tot[i][2 * j + 1]);
    }
  }
}

struct MlpPeBwdArgs {
  const float* dpe;      // [tiles*128, 96]
  const float* xyz;      // [rows,3]
  int64_t rows;
  int tpb, blocks;
  float* dL_dxyz;        // [rows,3] (summed over the time blocks) or NULL
  // not blender: tfeat = PE_10(t) -> per-row dL/dt straight from the pe cotangent (plain per-row call only)
  float* dL_dt_rows;     // [rows] or NULL
  const float* t; int64_t t_stride;
};

__global__ void __launch_bounds__(256) mlp_pe_bwd_kernel(MlpPeBwdArgs a) {
  const int64_t total = (int64_t)a.blocks * a.rows;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int blk = (int)(i / a.rows);
    const int64_t prow = i - (int64_t)blk * a.rows;
    const float* g = a.dpe + ((size_t)blk * a.tpb * kRows + prow) * 96;
    if (a.dL_dxyz) {
      float out[3];
#pragma unroll
      for (int d = 0; d < 3; d++) {
        const float x = a.xyz[3 * prow + d];
        float acc = g[d];
#pragma unroll
        for (int f = 0; f < 10; f++) {
          const float fr = (float)(1 << f);
          float sn, cs;
          sincosf(x * fr, &sn, &cs);
          acc += (g[3 + 6 * f + d] * cs - g[3 + 6 * f + 3 + d] * sn) * fr;
        }
        out[d] = acc;
      }
      if (a.blocks > 1) {
#pragma unroll
        for (int d = 0; d < 3; d++) atomicAdd(a.dL_dxyz + 3 * prow + d, out[d]);
      } else {
#pragma unroll
        for (int d = 0; d < 3; d++) a.dL_dxyz[3 * prow + d] = out[d];
      }
    }
    if (a.dL_dt_rows) {
      const float tv = a.t[prow * a.t_stride];
      float dt = g[63];
#pragma unroll
      for (int f = 0; f < 10; f++) {
        const float fr = (float)(1 << f);
        float sn, cs;
        sincosf(tv * fr, &sn, &cs);
        dt += (g[64 + 2 * f] * cs - g[65 + 2 * f] * sn) * fr;
      }
      a.dL_dt_rows[prow] = dt;
    }
  }
}

// Per-row timenet backward (blender, plain call whose rows carry their own time value; scene/deform_model.py:260-273
// accepts arbitrary [N,T], train_baseline.py:110-115 passes a materialised [N,1]).  Thread k owns hidden unit k and
// keeps its slices of the four parameter gradients in registers over the CTA's rows; rows are processed 32 at a time.
struct MlpTimenetRowsArgs {
  const float* dpe;        // [rows(+pad), 96]; columns 63..92 = d tfeat
  const float* t; int64_t t_stride;
  int64_t rows;
  const float* thidden;    // [rows,256], or [1,256] when the values turned out equal
  const float* tn0_w;      // [256,13]
  const float* tn2_w;      // [30,256]
  float* g_tn0_w; float* g_tn0_b; float* g_tn2_w; float* g_tn2_b;
  float* dL_dt_rows;       // [rows] or NULL
  const int* nonuniform;
};

__global__ void __launch_bounds__(256) mlp_timenet_rows_kernel(MlpTimenetRowsArgs a) {
  __shared__ float s_dtf[32][32];
  __shared__ float s_e[32][16];
  __shared__ float s_de[8][32][13];
  const int nonuni = *a.nonuniform;
  if (!nonuni && !a.dL_dt_rows) return;   // equal values and nobody wants per-row dL/dt: mlp_finish_kernel did it
  const int k = threadIdx.x, warp = k >> 5, lane = k & 31;
  float w2c[30], w0r[13], acc_w2[30], acc_w0[13], acc_b0 = 0.f;
#pragma unroll
  for (int i = 0; i < 30; i++) { w2c[i] = a.tn2_w[i * 256 + k]; acc_w2[i] = 0.f; }
#pragma unroll
  for (int c = 0; c < 13; c++) { w0r[c] = a.tn0_w[k * 13 + c]; acc_w0[c] = 0.f; }
  float acc_b2 = 0.f;   // threads 0..29: sum of d tfeat[k]
  const int64_t nbatch = (a.rows + 31) / 32;
  for (int64_t bt = blockIdx.x; bt < nbatch; bt += gridDim.x) {
    const int64_t r0 = bt * 32;
    __syncthreads();
    for (int idx = k; idx < 32 * 32; idx += 256) {
      const int rr = idx >> 5, j = idx & 31;
      const int64_t r = r0 + rr;
      s_dtf[rr][j] = (r < a.rows && j < 30) ? a.dpe[(size_t)r * 96 + 63 + j] : 0.f;
    }
    for (int idx = k; idx < 32 * 13; idx += 256) {
      const int rr = idx / 13, j = idx - rr * 13;
      const int64_t r = r0 + rr;
      float v = 0.f;
      if (r < a.rows) {
        const float tv = a.t[r * a.t_stride];
        if (j == 0) v = tv;
        else {
          const int f = (j - 1) >> 1;
          const float arg = tv * (float)(1 << f);
          v = ((j - 1) & 1) ? cosf(arg) : sinf(arg);
        }
      }
      s_e[rr][j] = v;
    }
    __syncthreads();
    for (int rr = 0; rr < 32; rr++) {
      const int64_t r = r0 + rr;
      if (r >= a.rows) break;
      const float hk = a.thidden[(nonuni ? (size_t)r * 256 : 0) + k];
      float dh = 0.f;
#pragma unroll
      for (int i = 0; i < 30; i++) {
        const float d = s_dtf[rr][i];
        acc_w2[i] = fmaf(d, hk, acc_w2[i]);
        dh = fmaf(w2c[i], d, dh);
      }
      if (!(hk > 0.f)) dh = 0.f;
      acc_b0 += dh;
#pragma unroll
      for (int c = 0; c < 13; c++) acc_w0[c] = fmaf(dh, s_e[rr][c], acc_w0[c]);
      if (a.dL_dt_rows) {
#pragma unroll
        for (int c = 0; c < 13; c++) {
          float v = dh * w0r[c];
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
          if (lane == 0) s_de[warp][rr][c] = v;
        }
      }
    }
    if (k < 30) {
#pragma unroll 4
      for (int rr = 0; rr < 32; rr++) acc_b2 += s_dtf[rr][k];
    }
    if (a.dL_dt_rows) {
      __syncthreads();
      if (k < 32 && r0 + k < a.rows) {
        float de[13];
#pragma unroll
        for (int c = 0; c < 13; c++) {
          float v = 0.f;
#pragma unroll
          for (int w = 0; w < 8; w++) v += s_de[w][k][c];
          de[c] = v;
        }
        float dt = de[0];
#pragma unroll
        for (int f = 0; f < 6; f++) dt += (de[1 + 2 * f] * s_e[k][2 + 2 * f] - de[2 + 2 * f] * s_e[k][1 + 2 * f]) * (float)(1 << f);
        a.dL_dt_rows[r0 + k] = dt;
      }
    }
  }
  if (!nonuni) return;   // equal values: the parameter gradients came from mlp_finish_kernel's shortcut
#pragma unroll
  for (int i = 0; i < 30; i++) atomicAdd(&a.g_tn2_w[i * 256 + k], acc_w2[i]);
  if (k < 30) atomicAdd(&a.g_tn2_b[k], acc_b2);
  atomicAdd(&a.g_tn0_b[k], acc_b0);
#pragma unroll
  for (int c = 0; c < 13; c++) atomicAdd(&a.g_tn0_w[k * 13 + c], acc_w0[c]);
}

}  // namespace d3g

using namespace d3g;

static int mlp_backward_impl(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p, int64_t n,
                             int blocks, const float* xyz, const float* t, int64_t t_stride, void* workspace,
                             const float* dL_dxyz_out, const float* dL_drot, const float* dL_dscale,
                             const d3g_mlp_grads* g, float* dL_dxyz, float* dL_dt, cudaStream_t stream) {
  if (int e = check_cfg(cfg)) return e;
  if (!packed || !p || !workspace || !g) return set_err(D3G_ERR_INVALID, "NULL argument");
  if (n < 0 || blocks < 1 || (n > 0 && (!xyz || !t))) return set_err(D3G_ERR_INVALID, "NULL tensor");
  for (int l = 0; l < 8; l++)
    if (!g->linear_w[l] || !g->linear_b[l]) return set_err(D3G_ERR_INVALID, "linear.%d gradient buffers missing", l);
  if (cfg->is_blender && (!g->timenet0_w || !g->timenet0_b || !g->timenet2_w || !g->timenet2_b))
    return set_err(D3G_ERR_INVALID, "timenet gradient buffers missing");
  if (n == 0) return 0;
  const MlpShape sh = mlp_shape(n, blocks);
  MlpWs ws = mlp_ws_view(workspace, n, true, blocks);
  const int tiles = sh.tiles;
  const int pe_dim = cfg->is_blender ? 93 : 84;
  const bool per_row = blocks == 1 && t_stride != 0 && n > 1;   // same rule as forward
  D3G_CUDA_TRY(cudaMemsetAsync(ws.scratch, 0, ws.scratch_floats * 4, stream));
  float* headb_acc = ws.scratch;
  float* db_fresh = ws.scratch + 16;
  int* nonuniform = ws.status + 1;
  uint32_t* amax_bits = (uint32_t*)(ws.scratch + 16 + (size_t)sh.blocks * 512);
  const int* gain_e = (const int*)((const uint8_t*)packed + kGainOff);
  const int gtarget = 5;   // largest scaled head cotangent in [32, 64)

  // ---- kernel 1: input-gradient chain ----
  {
    MlpBwdArgs a;
    a.wbwd = (const uint8_t*)packed + kBwdOff;
    a.rows = n; a.tiles = tiles; a.tpb = sh.tpb; a.is_6dof = cfg->is_6dof; a.heads = cfg->heads;
    a.mask = ws.mask; a.heads_out = ws.heads_out;
    a.g_xyz = dL_dxyz_out;
    a.g_rot = (cfg->heads & 2) ? dL_drot : nullptr;
    a.g_scale = (cfg->heads & 4) ? dL_dscale : nullptr;
    a.dz_img = ws.dz_img; a.dout_img = ws.dout_img; a.headb_acc = headb_acc; a.abort_flag = ws.status;
    a.amax_bits = amax_bits; a.gain_e = gain_e; a.gtarget = gtarget;
    { ProfScope _p(PROF_MLP_AMAX, stream);
      mlp_amax_kernel<<<2 * num_sms(), 256, 0, stream>>>(a);
      D3G_LAUNCH_CHECK(0, stream); }
    size_t smem = BwdSmem::total + 1024;
    D3G_CUDA_TRY(cudaFuncSetAttribute(mlp_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = tiles < mlp_grid_limit() ? tiles : mlp_grid_limit();
    { ProfScope _p(PROF_MLP_DGRAD, stream);
    mlp_dgrad_kernel<<<grid, kBwdThreads, smem, stream>>>(a);
    D3G_LAUNCH_CHECK(0, stream); }
  }
  // ---- kernel 2: weight gradients ----
  d3g_mlp_params gp;  // the gradient struct has the same layout; reuse head_map on it
  memcpy(&gp, g, sizeof(gp));
  HeadMap hm = head_map(*cfg, gp);
  {
    MlpWgradArgs a;
    memset(&a, 0, sizeof(a));
    { const char* e = getenv("D3G_WGRAD_DBG"); a.dbg = e ? atoi(e) : 0; }
    a.tiles = tiles; a.tpb = sh.tpb; a.abort_flag = ws.status;
    a.amax_bits = amax_bits; a.gain_e = gain_e; a.gtarget = gtarget;
    const size_t layer_stride = (size_t)tiles * kHBytes;
    int nit = 0;
    auto trunk = [&](int l, const uint8_t* b_img, uint32_t b_tile, int N, int bias, int col0, int ncols, int cost) {
      WgItem& w = a.items[nit++];
      w.a_img = ws.dz_img + (size_t)l * layer_stride;
      w.b_img = b_img; w.b_tile_bytes = b_tile; w.b_panels = (N + 63) / 64; w.N = N;
      w.bias = bias; w.kind = 0; w.layer = l; w.cost = cost;
      w.dW = g->linear_w[l]; w.ld = l == 0 ? pe_dim : (l == 5 ? pe_dim + 256 : 256); w.col0 = col0; w.ncols = ncols;
      w.db = bias ? g->linear_b[l] : nullptr;
      w.db_fresh = (bias && l == 0) ? db_fresh : ((bias && l == 5) ? db_fresh + 256 : nullptr);
    };
    for (int l = 7; l >= 1; l--)
      trunk(l, ws.h_img + (size_t)(l - 1) * layer_stride, kHBytes, 256, 1, l == 5 ? pe_dim : 0, 256, 128);
    trunk(0, ws.pe_img, kPeBytes, 96, 1, 0, pe_dim, 96);
    trunk(5, ws.pe_img, kPeBytes, 96, 0, 0, pe_dim, 96);
    {
      WgItem& w = a.items[nit++];
      w.a_img = ws.h_img + (size_t)7 * layer_stride;
      w.b_img = ws.dout_img; w.b_tile_bytes = kDoutBytes; w.b_panels = 1; w.N = 16;
      w.bias = 0; w.kind = 1; w.layer = -1; w.cost = 80;
    }
    a.n_items = nit;
    for (int j = 0; j < 16; j++) a.head_dw[j] = hm.w[j] ? (float*)hm.w[j] + (size_t)hm.r[j] * 256 : nullptr;
    size_t smem = (size_t)kWgStages * kWgStageBytes + 256 + 1024;
    D3G_CUDA_TRY(cudaFuncSetAttribute(mlp_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int grid = mlp_grid_limit();
    { ProfScope _p(PROF_MLP_WGRAD, stream);
    mlp_wgrad_kernel<<<grid, kWgThreads, smem, stream>>>(a);
    D3G_LAUNCH_CHECK(0, stream); }
  }
  // ---- kernel 3: head biases + shared-t timenet backward (one CTA per time block) ----
  {
    MlpFinishArgs f;
    memset(&f, 0, sizeof(f));
    f.headb_acc = headb_acc;
    for (int j = 0; j < 16; j++) f.head_db[j] = hm.b[j] ? (float*)hm.b[j] + hm.r[j] : nullptr;
    f.is_blender = cfg->is_blender; f.per_row = per_row ? 1 : 0; f.nonuniform = nonuniform;
    f.db_fresh = db_fresh; f.w0 = p->linear_w[0]; f.w5 = p->linear_w[5]; f.pe_dim = pe_dim;
    f.t = t; f.t_step = blocks > 1 ? 1 : 0; f.thidden = ws.thidden; f.tn0_w = p->timenet0_w; f.tn2_w = p->timenet2_w;
    f.g_tn0_w = g->timenet0_w; f.g_tn0_b = g->timenet0_b; f.g_tn2_w = g->timenet2_w; f.g_tn2_b = g->timenet2_b;
    f.dL_dt_blocks = per_row ? nullptr : dL_dt;
    { ProfScope _p(PROF_MLP_FINISH, stream);
    mlp_finish_kernel<<<blocks, 256, 0, stream>>>(f);
    D3G_LAUNCH_CHECK(0, stream); }
  }
  // ---- general input-block backward: dL/dxyz, per-row dL/dt, per-row timenet ----
  const bool want_dt_rows = per_row && dL_dt != nullptr;
  const bool rows_timenet = per_row && cfg->is_blender;
  if (dL_dxyz || want_dt_rows || rows_timenet) {
    MlpDpeArgs d;
    d.dz0 = ws.dz_img; d.dz5 = ws.dz_img + (size_t)5 * tiles * kHBytes;
    d.w0 = p->linear_w[0]; d.w5 = p->linear_w[5]; d.pe_dim = pe_dim; d.tiles = tiles; d.dpe = ws.dpe;
    d.skip_if_uniform = (!dL_dxyz && !want_dt_rows) ? 1 : 0; d.nonuniform = nonuniform;
    d.amax_bits = amax_bits; d.gain_e = gain_e; d.gtarget = gtarget;
    { ProfScope _p(PROF_MLP_DPE, stream);
    mlp_dpe_kernel<<<tiles < 4 * num_sms() ? tiles : 4 * num_sms(), 256, 0, stream>>>(d);
    D3G_LAUNCH_CHECK(0, stream); }
    const bool dt_from_pe = want_dt_rows && !cfg->is_blender;
    if (dL_dxyz || dt_from_pe) {
      if (dL_dxyz && blocks > 1) D3G_CUDA_TRY(cudaMemsetAsync(dL_dxyz, 0, (size_t)n * 3 * 4, stream));
      MlpPeBwdArgs e;
      e.dpe = ws.dpe; e.xyz = xyz; e.rows = n; e.tpb = sh.tpb; e.blocks = blocks; e.dL_dxyz = dL_dxyz;
      e.dL_dt_rows = dt_from_pe ? dL_dt : nullptr; e.t = t; e.t_stride = t_stride;
      const int64_t total = (int64_t)blocks * n;
      int64_t gb = (total + 255) / 256;
      if (gb > 16 * (int64_t)num_sms()) gb = 16 * (int64_t)num_sms();
      { ProfScope _p(PROF_MLP_DPE, stream);
      mlp_pe_bwd_kernel<<<(unsigned)gb, 256, 0, stream>>>(e);
      D3G_LAUNCH_CHECK(0, stream); }
    }
    if (rows_timenet) {
      MlpTimenetRowsArgs r;
      r.dpe = ws.dpe; r.t = t; r.t_stride = t_stride; r.rows = n; r.thidden = ws.thidden;
      r.tn0_w = p->timenet0_w; r.tn2_w = p->timenet2_w;
      r.g_tn0_w = g->timenet0_w; r.g_tn0_b = g->timenet0_b; r.g_tn2_w = g->timenet2_w; r.g_tn2_b = g->timenet2_b;
      r.dL_dt_rows = dL_dt; r.nonuniform = nonuniform;
      int64_t gb = (n + 31) / 32;
      if (gb > 2 * (int64_t)num_sms()) gb = 2 * (int64_t)num_sms();
      { ProfScope _p(PROF_MLP_DPE, stream);
      mlp_timenet_rows_kernel<<<(unsigned)gb, 256, 0, stream>>>(r);
      D3G_LAUNCH_CHECK(0, stream); }
    }
  }
  return 0;
}

extern "C" int d3g_mlp_backward(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p,
                                int64_t rows, const float* xyz, const float* t, int64_t t_stride, void* workspace,
                                const float* dL_dxyz_out, const float* dL_drot, const float* dL_dscale,
                                const d3g_mlp_grads* g, float* dL_dxyz, float* dL_dt, void* stream_) {
  return mlp_backward_impl(cfg, packed, p, rows, 1, xyz, t, t_stride, workspace, dL_dxyz_out, dL_drot, dL_dscale, g,
                           dL_dxyz, dL_dt, (cudaStream_t)stream_);
}

extern "C" int d3g_mlp_backward_times(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p,
                                      int64_t n_points, int32_t n_times, const float* xyz, const float* t_values,
                                      void* workspace, const float* dL_dxyz_out, const float* dL_drot,
                                      const float* dL_dscale, const d3g_mlp_grads* g, float* dL_dxyz, float* dL_dt,
                                      void* stream_) {
  if (n_times < 1) return set_err(D3G_ERR_INVALID, "n_times must be >= 1");
  return mlp_backward_impl(cfg, packed, p, n_points, n_times, xyz, t_values, 0, workspace, dL_dxyz_out, dL_drot,
                           dL_dscale, g, dL_dxyz, dL_dt, (cudaStream_t)stream_);
}

// Parity / debug access: activation pattern of trunk layer `layer` as saved by a forward with
// save_for_backward = 1.  dst_host[row*8 + c] bit j = (h_layer[row][c*32+j] > 0).  Synchronous.
extern "C" int d3g_mlp_read_masks(const d3g_mlp_config* cfg, const void* workspace, int64_t rows, int32_t layer,
                                  uint32_t* dst_host, void* stream_) {
  if (int e = check_cfg(cfg)) return e;
  if (!workspace || !dst_host || rows <= 0 || layer < 0 || layer > 7) return set_err(D3G_ERR_INVALID, "bad arguments");
  cudaStream_t stream = (cudaStream_t)stream_;
  MlpWs ws = mlp_ws_view((void*)workspace, rows, true);
  const size_t tiles = (size_t)((rows + kRows - 1) / kRows);
  uint32_t* tmp = (uint32_t*)malloc(tiles * 8 * 128 * 4);
  if (!tmp) return set_err(D3G_ERR_CAPACITY, "out of host memory");
  cudaError_t e = cudaMemcpyAsync(tmp, ws.mask + (size_t)layer * tiles * 8 * 128, tiles * 8 * 128 * 4,
                                  cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) { free(tmp); return set_err(D3G_ERR_CUDA, "mask read failed: %s", cudaGetErrorString(e)); }
  for (int64_t r = 0; r < rows; r++)
    for (int c = 0; c < 8; c++) dst_host[r * 8 + c] = tmp[((size_t)(r / kRows) * 8 + c) * 128 + (r % kRows)];
  free(tmp);
  return 0;
}
