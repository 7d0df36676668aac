// Fused deformation-MLP forward for sm_100a (tcgen05 / TMEM / bulk-copy).
//
// Replaces DeformNetwork(.Baseline).forward, utils/time_utils.py:102-127 / :176-201: positional
// encoding, 8x256 trunk with the skip at layer 5, heads and the optional se(3) exponential map run in
// ONE persistent kernel.  Arithmetic is the split-fp16 scheme of mlp_common.cuh (three kind::f16 products per
// k-step, fp32 accumulation: fp32-class accuracy).  Per 128-point tile:
//   * the positional-encoding block (96 columns, hi | lo) is built in shared memory once and feeds layer 0 and
//     the skip layer as an smem A operand;
//   * every other activation lives ONLY in tensor memory: the fp32 accumulator of layer l (TMEM half
//     l&1, 256 columns) is bias+ReLU'd by the epilogue warps and written back IN PLACE as packed fp16 — the
//     32 accumulator columns of a chunk become 16 columns of hi pairs + 16 columns of lo pairs — and is then
//     consumed directly as the TMEM A operand of layer l+1, whose accumulator is the other half.
//     Activations never touch shared memory or HBM in inference;
//   * weights stream from the packed image (L2-resident, 2 MB) through a bulk-copy ring
//     (cp.async.bulk + mbarrier expect_tx), 32 KB = 256 outputs x (32 k hi | 32 k lo) per stage;
//   * the epilogue hands the next layer its A operand 32 features at a time (8 mbarriers), so the MMA
//     of layer l+1 starts while the epilogue of layer l is still draining.
// Warp roles (kEpiWarps = 8): 0-7 epilogue/PE (warp w owns TMEM lanes 32*(w&3).., column chunks of parity w>>2),
// 8 weight producer, 9 MMA issuer + TMEM owner.
#include <math.h>
#include <stdlib.h>

#include "mlp_common.cuh"

namespace d3g {

constexpr int kNumStages = 4;
constexpr int kEpiWarps = 8;                        // 2 per TMEM lane quarter, 4 column chunks each per layer
constexpr int kFwdThreads = (kEpiWarps + 2) * 32;   // + weight producer + MMA issuer

struct FwdSmem {
  // offsets inside dynamic shared memory (after 1024-alignment)
  static constexpr uint32_t pe = 0;
  static constexpr uint32_t stages = kPeSmemBytes;
  static constexpr uint32_t store = stages + kNumStages * kStageBytes;   // [4 lane quarters][2] x 4 KB bulk-store staging
  static constexpr uint32_t bias = store + 8 * 4096;
  static constexpr uint32_t bars = bias + kBiasFloats * 4 + 64;  // 8-byte aligned
  static constexpr uint32_t total = bars + 320;
};
static_assert(FwdSmem::bars % 8 == 0, "barrier alignment");

struct MlpFwdArgs {
  const uint8_t* wfwd;
  const float* bias;
  const float* xyz;
  int64_t rows;          // points per time block
  int tiles;             // all tiles of the launch (blocks * tpb)
  int tpb;               // tiles per time block
  const float* tfeat;
  int per_row_t;         // rows carry their own time value (plain call with t_stride != 0): the stride is 32 unless
  const int* nonuniform; // ... the device found all values equal (*nonuniform == 0), then row 0 serves every row
  int pe_dim, t_dim;
  int is_6dof, heads;
  int save;
  uint8_t* pe_img;
  uint8_t* h_img;
  uint32_t* mask;
  float* heads_out;
  float* d_xyz;
  float* d_rot;
  float* d_scale;
  int* abort_flag;
  int dbg;   // timing experiments only (D3G_FWD_DBG): bit0 skip tcgen05.st, bit1 skip tcgen05.ld, bit3 1/8 weight bytes,
             // bit4 one product instead of three, bit5 clock stamps of CTA 0's second tile into status[16..]
  int* tl;   // timeline buffer (status + 16), 224 ints
};
#define TL(slot)                                                                                     \
  do {                                                                                               \
    if ((a.dbg & 32) && blockIdx.x == 0 && tile == (int)gridDim.x) a.tl[slot] = (int)(clock64() - tl0); \
  } while (0)

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void stg128(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(kFwdThreads, 1) mlp_fwd_kernel(MlpFwdArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sPE = smem + FwdSmem::pe;
  uint8_t* sStage = smem + FwdSmem::stages;
  float* sBias = (float*)(smem + FwdSmem::bias);
  uint8_t* sStore = smem + FwdSmem::store;
  uint64_t* bars = (uint64_t*)(smem + FwdSmem::bars);
  uint64_t* full = bars;                   // [kNumStages]
  uint64_t* empty = bars + kNumStages;     // [kNumStages]
  uint64_t* acc_full = bars + 2 * kNumStages;
  uint64_t* pe_ready = acc_full + 1;
  // go[c]: "chunk c (32 features) of the next layer may be multiplied" = its A operand has been handed over by the four
  // epilogue warps that produce it AND its weight stage has landed (the producer's expect_tx lands on the same
  // barrier).  One wait per chunk on the issuing thread instead of two: an mbarrier poll costs that thread ~200 clk
  // and it shares its scheduler with two epilogue warps (measured 1250 clk per chunk against 810 clk of tensor work
  // while the epilogue runs).
  uint64_t* go = pe_ready + 1;             // [8]
  // The heads step has its own hand-over barriers (8 epilogue warps only; its single weight stage is waited for
  // through full[s]): the producer runs up to kNumStages stages ahead, i.e. into the NEXT tile's layer-1 stages while
  // layer 7's epilogue is still handing over, so those arrivals must not meet the heads phase on go[].
  uint64_t* goh = go + 8;                  // [8]
  uint32_t* tmem_slot = (uint32_t*)(goh + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  volatile int* abort_flag = a.abort_flag;

  if (tid == 0) {
    for (int i = 0; i < kNumStages; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(acc_full, 1);
    mbar_init(pe_ready, 8);
    for (int i = 0; i < 8; i++) { mbar_init(&go[i], 5); mbar_init(&goh[i], 4); }   // 4 epilogue warps (+ 1 weight stage)
    fence_barrier_init();
  }
  if (warp == kEpiWarps + 1) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < kBiasFloats; i += kFwdThreads) sBias[i] = a.bias[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const long long tl0 = clock64();

  if (warp == kEpiWarps) {
    // ================= weight producer =================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      bool ok = true;
      for (int tile = blockIdx.x; ok && tile < a.tiles; tile += gridDim.x) {
        for (int st = 0; st < kFwdStages; st++) {
          if (!mbar_wait(&empty[s], ph ^ 1, abort_flag)) { ok = false; break; }
          int layer, kp;
          fwd_stage_info(st, layer, kp);
          uint64_t* bar = &full[s];                       // pe-fed stages (layer 0, pe part of layer 5)
          uint32_t bytes = kStageBytes;
          if (layer == 9) {                               // heads: one 16 KB stage, waited for through full[s]
            bytes = 16384u;
          } else if (layer != 0 && layer != 8) {
            bar = &go[kp];
          }
          mbar_arrive_expect_tx(bar, bytes);
          bulk_g2s(sStage + (size_t)s * kStageBytes, a.wfwd + (size_t)st * kStageBytes, bytes, bar);
          if (++s == kNumStages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == kEpiWarps + 1) {
    // ================= MMA issuer =================
    // One k-chunk = 32 features: the B rows hold [32 k of Whi | 32 k of Wlo] (4 descriptors 32 B apart), the A
    // side [hi | lo] likewise (16 + 16 packed TMEM columns, or 64 + 64 bytes of a shared-memory row).
    if (lane == 0) {
      const uint32_t idesc256 = make_idesc_f16(128, 256, 0, 0);
      const uint32_t idesc16 = make_idesc_f16(128, 16, 0, 0);
      int s = 0;
      uint32_t ph = 0, pe_ph = 0, a_ph = 0, h_ph = 0;
      bool ok = true;
      // (this thread shares its scheduler with two epilogue warps: everything it executes per MMA counts — descriptors
      // are advanced from two precomputed bases, barrier waits are non-suspending polls)
      const uint64_t bdesc0 = make_smem_desc(smem_u32(sStage), 16, 1024), adesc0 = make_smem_desc(smem_u32(sPE), 16, 1024);
      // full[s] only completes for the pe-fed stages that pass through slot s: its parity is tracked per slot
      uint32_t full_bits = 0;
      auto wait_full = [&]() {
        const bool r = mbar_wait(&full[s], (full_bits >> s) & 1u, abort_flag);
        full_bits ^= 1u << s;
        return r;
      };
      auto release = [&]() { umma_commit(&empty[s]); if (++s == kNumStages) { s = 0; ph ^= 1; } };
      for (int tile = blockIdx.x; ok && tile < a.tiles; tile += gridDim.x) {
        if (!mbar_wait(pe_ready, pe_ph, abort_flag)) { ok = false; break; }
        pe_ph ^= 1;
        tc_fence_after();
        // pe-fed part of a layer: A from shared memory (3 chunks)
        auto pe_part = [&](uint32_t d_col) {
          for (int kc = 0; ok && kc < 3; kc++) {
            if (!wait_full()) { ok = false; break; }
            tc_fence_after();
            const uint64_t sb = desc_adv(bdesc0, (uint32_t)s * kStageBytes);
            const uint64_t sa = desc_adv(adesc0, (uint32_t)kc * kPanelBytes);
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {   // hi * Whi
              umma_f16_ss(tmem_base + d_col, desc_adv(sa, ks * 32), desc_adv(sb, ks * 32), idesc256, (kc | ks) != 0);
            }
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {   // lo * Whi, hi * Wlo
              umma_f16_ss(tmem_base + d_col, desc_adv(sa, 64 + ks * 32), desc_adv(sb, ks * 32), idesc256, 1u);
              umma_f16_ss(tmem_base + d_col, desc_adv(sa, ks * 32), desc_adv(sb, 64 + ks * 32), idesc256, 1u);
            }
            release();
          }
        };
        // h-fed part: A from tensor memory, handed over 32 features at a time
        auto h_part = [&](uint32_t d_col, uint32_t a_col, bool accumulate_first) {
          for (int kc = 0; ok && kc < 8; kc++) {
            if (!mbar_wait(&go[kc], a_ph, abort_flag)) { ok = false; break; }
            tc_fence_after();
            if (d_col == 256 && a_col == 0) TL(16 + kc);          // odd layers: chunk kc may be multiplied
            const uint64_t sb = desc_adv(bdesc0, (uint32_t)s * kStageBytes);
            const uint32_t ta = tmem_base + a_col + kc * 32;
#pragma unroll
            for (int ks = 0; ks < 2; ks++)
              umma_f16_ts(tmem_base + d_col, ta + ks * 8, desc_adv(sb, ks * 32), idesc256,
                          (accumulate_first || (kc | ks) != 0) ? 1u : 0u);
            if (!(a.dbg & 16)) {
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
              umma_f16_ts(tmem_base + d_col, ta + 16 + ks * 8, desc_adv(sb, ks * 32), idesc256, 1u);
              umma_f16_ts(tmem_base + d_col, ta + ks * 8, desc_adv(sb, 64 + ks * 32), idesc256, 1u);
            }
            }
            release();
          }
          a_ph ^= 1;
        };
        TL(0);
        pe_part(0);                       // layer 0 -> half 0
        if (!ok) break;
        umma_commit(acc_full);
        TL(1);
        for (int l = 1; ok && l < 8; l++) {
          uint32_t d_col = (l & 1) * 256, a_col = ((l - 1) & 1) * 256;
          if (l == 5) {
            pe_part(d_col);               // overlaps the epilogue of layer 4
            if (!ok) break;
            h_part(d_col, a_col, true);
          } else {
            h_part(d_col, a_col, false);
          }
          if (!ok) break;
          umma_commit(acc_full);
          TL(1 + l);
        }
        if (!ok) break;
        // heads: N = 16, A = h7 (half 1), D = half 0 columns 0..15, single 16 KB stage of 8 chunks x 2 KB
        {
          if (!wait_full()) { ok = false; break; }
          for (int kc = 0; ok && kc < 8; kc++) {
            if (!mbar_wait(&goh[kc], h_ph, abort_flag)) { ok = false; break; }
            tc_fence_after();
            const uint64_t sb = desc_adv(bdesc0, (uint32_t)s * kStageBytes + (uint32_t)kc * 2048u);
            const uint32_t ta = tmem_base + 256 + kc * 32;
#pragma unroll
            for (int ks = 0; ks < 2; ks++)
              umma_f16_ts(tmem_base + 0, ta + ks * 8, desc_adv(sb, ks * 32), idesc16, (kc | ks) != 0);
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
              umma_f16_ts(tmem_base + 0, ta + 16 + ks * 8, desc_adv(sb, ks * 32), idesc16, 1u);
              umma_f16_ts(tmem_base + 0, ta + ks * 8, desc_adv(sb, 64 + ks * 32), idesc16, 1u);
            }
          }
          if (!ok) break;
          h_ph ^= 1;
          release();
          umma_commit(acc_full);
        }
      }
    }
  } else {
    // ================= PE + epilogue warps (256 threads) =================
    const int q = warp & 3, cg = warp >> 2, chalf = cg;   // cg: column-chunk parity (PE: column half)
    const int row = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t acc_ph = 0;
    bool ok = true;
    const int64_t tfeat_stride = (a.per_row_t && *a.nonuniform) ? 32 : 0;
    for (int tile = blockIdx.x; ok && tile < a.tiles; tile += gridDim.x) {
      const int blk = tile / a.tpb;
      const int64_t prow = (int64_t)(tile - blk * a.tpb) * kRows + row;   // point index inside the time block
      const bool live = prow < a.rows;
      const int64_t grow = (int64_t)blk * a.rows + prow;                  // row of the [blocks, rows, .] outputs
      // ---- positional encoding block: each row is built by two threads (columns 0..47 / 48..95) in
      // registers, split into fp16 hi / lo and written to shared memory (three k-chunks of [32 hi | 32 lo]
      // K-major SW128 rows: A operand of layer 0 and of the skip layer) and, when saving for backward, to HBM
      // as the hi-part [point][feature] image the weight gradient reads ----
      {
        float x[3] = {0.f, 0.f, 0.f};
        if (live) { x[0] = a.xyz[3 * prow]; x[1] = a.xyz[3 * prow + 1]; x[2] = a.xyz[3 * prow + 2]; }
        float pe[48];
        if (chalf == 0) {
          pe[0] = x[0]; pe[1] = x[1]; pe[2] = x[2];
#pragma unroll
          for (int f = 0; f < 8; f++) {
            float fr = (float)(1 << f);
#pragma unroll
            for (int d = 0; d < 3; d++) {
              float sn, cs;
              sincosf(x[d] * fr, &sn, &cs);
              pe[3 + 6 * f + d] = sn;
              if (f < 7) pe[3 + 6 * f + 3 + d] = cs;
            }
          }
        } else {
#pragma unroll
          for (int f = 7; f < 10; f++) {
            float fr = (float)(1 << f);
#pragma unroll
            for (int d = 0; d < 3; d++) {
              float sn, cs;
              sincosf(x[d] * fr, &sn, &cs);
              if (f > 7) pe[6 * (f - 8) + 3 + d] = sn;
              pe[6 * (f - 8) + 6 + d] = cs;
            }
          }
          const float* tf = a.tfeat + (size_t)blk * 32 + (live ? prow * tfeat_stride : 0);
#pragma unroll
          for (int j = 0; j < 33; j++) pe[15 + j] = (live && j < a.t_dim) ? tf[j] : 0.f;
        }
        uint32_t hi[24], lo[24];
#pragma unroll
        for (int j = 0; j < 24; j++) split2(pe[2 * j], pe[2 * j + 1], hi[j], lo[j]);
#pragma unroll
        for (int g = 0; g < 6; g++) {   // groups of 8 columns = one 16-byte chunk of hi and one of lo
          const uint32_t col0 = chalf * 48 + 8 * g;
          const uint32_t base = smem_u32(sPE) + (col0 >> 5) * kPanelBytes + row * 128;
          const uint32_t ch = (col0 & 31u) >> 3;
          sts128(base + (((ch) ^ (row & 7)) << 4), hi[4 * g], hi[4 * g + 1], hi[4 * g + 2], hi[4 * g + 3]);
          sts128(base + (((4 + ch) ^ (row & 7)) << 4), lo[4 * g], lo[4 * g + 1], lo[4 * g + 2], lo[4 * g + 3]);
        }
        if (a.save) {
          uint8_t* dst = a.pe_img + (size_t)tile * kPeBytes + row * 128;
#pragma unroll
          for (int g = 0; g < 6; g++) {
            const uint32_t col0 = chalf * 48 + 8 * g;
            stg128(dst + (col0 >> 6) * kPanelBytes + ((((col0 & 63u) >> 3) ^ (row & 7)) << 4), hi[4 * g], hi[4 * g + 1],
                   hi[4 * g + 2], hi[4 * g + 3]);
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(pe_ready);

      // ---- trunk epilogues: bias + ReLU + hi/lo split in place in TMEM, 32 features at a time ----
      for (int l = 0; ok && l < 8; l++) {
        {   // one lane polls, the warp follows: 8 pollers instead of 256 on the barrier word
          int got = 1;
          if (lane == 0) got = mbar_wait(acc_full, acc_ph, abort_flag) ? 1 : 0;
          got = __shfl_sync(0xffffffffu, got, 0);
          if (!got) { ok = false; break; }
        }
        acc_ph ^= 1;
        tc_fence_after();
        if (lane == 0 && q == 0 && l < 3) TL(40 + l * 24 + cg * 12);
        const uint32_t half = (l & 1) * 256;
        const float* bl = sBias + l * 256;
        for (int c = cg; c < 8; c += kEpiWarps / 4) {
          uint32_t v[32];
          if (!(a.dbg & 2)) {
            tmem_ld32(lane_addr + half + c * 32, v);
            tmem_wait_ld();
          } else {
#pragma unroll
            for (int j = 0; j < 32; j++) v[j] = 0x3f000000u + j;
          }
          if (lane == 0 && q == 0 && l < 3) TL(40 + l * 24 + cg * 12 + 1 + (c >> 1) * 2);
          uint32_t bits = 0;
          float bv[32];
          const uint32_t bl_s = smem_u32(bl + c * 32);
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const float4 b4 = lds128f(bl_s + 16 * j);
            bv[4 * j] = b4.x; bv[4 * j + 1] = b4.y; bv[4 * j + 2] = b4.z; bv[4 * j + 3] = b4.w;
          }
          // z = acc * 2^-s + bias; `neg` is its sign smeared over the word: activation = z & ~neg, mask bit
          // ~neg & (1 << j) (z == +0 counts as active with value 0); then the hi/lo split, two features per word
          uint32_t o[32];
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float r[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
              const uint32_t zb = __float_as_uint(fmaf(__uint_as_float(v[j + e]), kWScaleInv, bv[j + e]));
              uint32_t neg, rb;
              asm("shr.s32 %0, %1, 31;" : "=r"(neg) : "r"(zb));
              asm("lop3.b32 %0, %1, %2, %2, 0x30;" : "=r"(rb) : "r"(zb), "r"(neg));                      // z & ~neg
              asm("lop3.b32 %0, %0, %1, %2, 0xF2;" : "+r"(bits) : "r"(neg), "r"(1u << (j + e)));          // bits | (~neg & bit)
              r[e] = __uint_as_float(rb);
            }
            split2(r[0], r[1], o[j >> 1], o[16 + (j >> 1)]);
          }
          if (!(a.dbg & 1)) tmem_st32(lane_addr + half + c * 32, o);
          // hand the chunk to the MMA warp first, then save it for backward
          tmem_wait_st();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(l == 7 ? &goh[c] : &go[c]);
          if (lane == 0 && q == 0 && l < 3) TL(40 + l * 24 + cg * 12 + 1 + (c >> 1) * 2 + 1);
          if (a.save) {
            // hi part -> [point][feature] image.  The two warps of a lane quarter hold the two 64-byte halves of the
            // same 32 rows of panel c/2 = one contiguous 4 KB block of the image: they assemble it in shared memory and
            // the bulk-copy engine writes it as one store (per-thread 16-byte global stores at a 128-byte lane stride
            // stalled the epilogue: 8.5k instead of 3.3k clk per layer, profiles/r2_mlp_fwd_timeline.txt).
            const int pr = c >> 1;
            uint8_t* stg = sStore + ((q * 2 + (pr & 1)) << 12);
            if (cg == 0 && lane == 0) bulk_wait_read<1>();     // the store issued from this buffer two pairs ago has left it
            named_bar_sync(1 + q, 64);
            const uint32_t line = smem_u32(stg) + lane * 128;
#pragma unroll
            for (int g = 0; g < 4; g++)
              sts128(line + (((4 * (c & 1) + g) ^ (lane & 7)) << 4), o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
            fence_proxy_async();
            named_bar_sync(1 + q, 64);
            if (cg == 0 && lane == 0) {
              bulk_s2g(a.h_img + ((size_t)l * a.tiles + tile) * kHBytes + (size_t)pr * kPanelBytes + (size_t)q * 4096, stg, 4096);
              bulk_commit();
            }
            a.mask[(((size_t)l * a.tiles + tile) * 8 + c) * 128 + row] = bits;
          }
        }
      }
      if (!ok) break;
      // ---- heads ----
      if (!mbar_wait(acc_full, acc_ph, abort_flag)) { ok = false; break; }
      acc_ph ^= 1;
      tc_fence_after();
      if (cg == 0) {
        uint32_t v[16];
        tmem_ld16(lane_addr + 0, v);
        tmem_wait_ld();
        float o[16];
        const float* bh = sBias + 8 * 256;
#pragma unroll
        for (int j = 0; j < 16; j++) o[j] = fmaf(__uint_as_float(v[j]), kWScaleInv, bh[j]);
        if (a.save) {
          float4* dst = (float4*)(a.heads_out + ((size_t)tile * kRows + row) * 16);
#pragma unroll
          for (int j = 0; j < 4; j++) dst[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        }
        if (live) {
          if (!a.is_6dof) {
            a.d_xyz[3 * grow] = o[0]; a.d_xyz[3 * grow + 1] = o[1]; a.d_xyz[3 * grow + 2] = o[2];
            if (a.d_rot) { *(float4*)(a.d_rot + 4 * grow) = make_float4(o[3], o[4], o[5], o[6]); }
            if (a.d_scale) { a.d_scale[3 * grow] = o[7]; a.d_scale[3 * grow + 1] = o[8]; a.d_scale[3 * grow + 2] = o[9]; }
          } else {
            // utils/time_utils.py:114-121 + utils/rigid_utils.py:60-83 (exp_se3)
            float w0 = o[0], w1 = o[1], w2 = o[2], v0 = o[3], v1 = o[4], v2 = o[5];
            float theta = sqrtf(w0 * w0 + w1 * w1 + w2 * w2);
            w0 = w0 / theta + 1e-5f; w1 = w1 / theta + 1e-5f; w2 = w2 / theta + 1e-5f;
            v0 = v0 / theta + 1e-5f; v1 = v1 / theta + 1e-5f; v2 = v2 / theta + 1e-5f;
            float sn, cs;
            sincosf(theta, &sn, &cs);
            float W[9] = {0.f, -w2, w1, w2, 0.f, -w0, -w1, w0, 0.f};
            float W2[9];
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
              for (int j = 0; j < 3; j++)
                W2[3 * i + j] = W[3 * i] * W[j] + W[3 * i + 1] * W[3 + j] + W[3 * i + 2] * W[6 + j];
            float R[9], Pm[9];
#pragma unroll
            for (int i = 0; i < 9; i++) {
              float id = (i == 0 || i == 4 || i == 8) ? 1.f : 0.f;
              R[i] = id + sn * W[i] + (1.0f - cs) * W2[i];
              Pm[i] = theta * id + (1.0f - cs) * W[i] + (theta - sn) * W2[i];
            }
            float p0 = Pm[0] * v0 + Pm[1] * v1 + Pm[2] * v2;
            float p1 = Pm[3] * v0 + Pm[4] * v1 + Pm[5] * v2;
            float p2 = Pm[6] * v0 + Pm[7] * v1 + Pm[8] * v2;
            float4* out = (float4*)(a.d_xyz + 16 * grow);
            out[0] = make_float4(R[0], R[1], R[2], p0);
            out[1] = make_float4(R[3], R[4], R[5], p1);
            out[2] = make_float4(R[6], R[7], R[8], p2);
            out[3] = make_float4(0.f, 0.f, 0.f, 1.f);
            if (a.d_rot) { *(float4*)(a.d_rot + 4 * grow) = make_float4(o[6], o[7], o[8], o[9]); }
            if (a.d_scale) { a.d_scale[3 * grow] = o[10]; a.d_scale[3 * grow + 1] = o[11]; a.d_scale[3 * grow + 2] = o[12]; }
          }
        }
      }
      tc_fence_before();
    }
    if (cg == 0 && lane == 0) bulk_wait_all<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps + 1) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------
// small helper kernels
// ------------------------------------------------------------------------------------------------

// Do the rows of a per-row time tensor carry different values?  (train_baseline.py:110-115 materialises
// `fid.expand(N,1) + ast_noise`, which for blender scenes is N copies of one value.)  flag[0] is zeroed by the
// caller and set to 1 when any element differs from t[0]; the kernels downstream branch on it, so the shared-t
// fast path needs no host round-trip.
__global__ void __launch_bounds__(256)
t_nonuniform_kernel(int64_t rows, const float* __restrict__ t, int64_t t_stride, int* __restrict__ flag) {
  const float t0 = t[0];
  int differs = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += (int64_t)gridDim.x * blockDim.x)
    differs |= (t[i * t_stride] != t0);
  if (__syncthreads_or(differs) && threadIdx.x == 0) *flag = 1;
}

// Time feature: blender -> timenet(PE_6(t)) = 13 -> 256 -> ReLU -> 30 (utils/time_utils.py:74-76,105);
// otherwise PE_10(t) (21 values).  One CTA of 256 threads per distinct time value (grid-stride).  With
// `nonuniform` given and clear, only row 0 is evaluated.
__global__ void __launch_bounds__(256)
tfeat_kernel(int64_t rows_t, const float* __restrict__ t, int64_t t_stride, int is_blender,
             const float* __restrict__ w0, const float* __restrict__ b0, const float* __restrict__ w2,
             const float* __restrict__ b2, float* __restrict__ tfeat, float* __restrict__ thidden,
             const int* __restrict__ nonuniform) {
  __shared__ float e[32];
  __shared__ float hid[256];
  if (nonuniform && !*nonuniform) rows_t = 1;
  for (int64_t r = blockIdx.x; r < rows_t; r += gridDim.x) {
    __syncthreads();
    float tv = t[r * t_stride];
    int L = is_blender ? 6 : 10;
    if (threadIdx.x < 32) {
      int j = threadIdx.x;
      float v = 0.f;
      if (j == 0) v = tv;
      else if (j < 1 + 2 * L) {
        int f = (j - 1) >> 1;
        float arg = tv * (float)(1 << f);
        v = ((j - 1) & 1) ? cosf(arg) : sinf(arg);
      }
      e[j] = v;
    }
    __syncthreads();
    if (!is_blender) {
      if (threadIdx.x < 32) tfeat[r * 32 + threadIdx.x] = threadIdx.x < 21 ? e[threadIdx.x] : 0.f;
      continue;
    }
    {
      int j = threadIdx.x;
      float acc = b0[j];
#pragma unroll
      for (int k = 0; k < 13; k++) acc += w0[j * 13 + k] * e[k];
      acc = fmaxf(acc, 0.f);
      hid[j] = acc;
      if (thidden) thidden[r * 256 + j] = acc;
    }
    __syncthreads();
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = warp; i < 32; i += 8) {
      float acc = 0.f;
      if (i < 30)
        for (int k = lane; k < 256; k += 32) acc += w2[i * 256 + k] * hid[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) tfeat[r * 32 + i] = i < 30 ? acc + b2[i] : 0.f;
    }
  }
}

// Pack parameters into the streaming images (forward stages, biases, transposed backward stages).  A stage is
// 256 rows x 128 B: bytes 0..63 of a row = fp16 hi parts of 32 consecutive k of W * 2^kWScaleLog2, bytes 64..127 =
// the fp16 lo parts (W*2^s - hi), 16-byte chunks XOR-swizzled with (row & 7).  One thread per (stage, row, 8 k).
__global__ void __launch_bounds__(256)
pack_weights_kernel(d3g_mlp_config cfg, d3g_mlp_params p, uint8_t* __restrict__ packed) {
  const int pe_dim = cfg.is_blender ? 93 : 84;
  int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int per_stage = 256 * 4;  // (row, group of 8 k)
  const int n_fwd = kFwdStages * per_stage, n_bwd = kBwdStages * per_stage;
  auto head_src = [&](int h, const float*& w, const float*& b, int& r) {
    w = nullptr; b = nullptr; r = 0;
    if (!cfg.is_6dof) {
      if (h < 3) { w = p.warp_w; b = p.warp_b; r = h; }
      else if (h < 7) { if (cfg.heads & 2) { w = p.rot_w; b = p.rot_b; r = h - 3; } }
      else if (h < 10) { if (cfg.heads & 4) { w = p.scale_w; b = p.scale_b; r = h - 7; } }
    } else {
      if (h < 3) { w = p.branch_w_w; b = p.branch_w_b; r = h; }
      else if (h < 6) { w = p.branch_v_w; b = p.branch_v_b; r = h - 3; }
      else if (h < 10) { if (cfg.heads & 2) { w = p.rot_w; b = p.rot_b; r = h - 6; } }
      else if (h < 13) { if (cfg.heads & 4) { w = p.scale_w; b = p.scale_b; r = h - 10; } }
    }
  };
  // write 8 values as one hi chunk (j) and one lo chunk (4 + j) of row n at `row_base`
  auto put = [&](uint8_t* row_base, int n, int j, const float (&v)[8]) {
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; e++) split2(v[2 * e] * kWScale, v[2 * e + 1] * kWScale, hi[e], lo[e]);
    *(uint4*)(row_base + ((j ^ (n & 7)) << 4)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *(uint4*)(row_base + (((4 + j) ^ (n & 7)) << 4)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  };
  if (gid < n_fwd) {
    int st = gid / per_stage, rem = gid % per_stage, n = rem >> 2, j = rem & 3;
    int layer, kp;
    fwd_stage_info(st, layer, kp);
    float v[8];
    if (layer == 9) {
      if (n >= 16) return;
      // heads stage: 8 k-chunks of 16 rows inside one 16 KB stage; this thread handles all 8 chunks
      const float* w; const float* b; int r;
      head_src(n, w, b, r);
      for (int kp2 = 0; kp2 < 8; kp2++) {
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = w ? w[r * 256 + kp2 * 32 + j * 8 + e] : 0.f;
        put(packed + (size_t)st * kStageBytes + kp2 * 2048 + n * 128, n, j, v);
      }
      return;
    }
    const float* W; int ld, col0, ncols;
    if (layer == 0) { W = p.linear_w[0]; ld = pe_dim; col0 = 0; ncols = pe_dim; }
    else if (layer == 8) { W = p.linear_w[5]; ld = pe_dim + 256; col0 = 0; ncols = pe_dim; }
    else if (layer == 5) { W = p.linear_w[5]; ld = pe_dim + 256; col0 = pe_dim; ncols = 256; }
    else { W = p.linear_w[layer]; ld = 256; col0 = 0; ncols = 256; }
#pragma unroll
    for (int e = 0; e < 8; e++) {
      int c = kp * 32 + j * 8 + e;
      v[e] = c < ncols ? W[(size_t)n * ld + col0 + c] : 0.f;
    }
    put(packed + (size_t)st * kStageBytes + n * 128, n, j, v);
    return;
  }
  gid -= n_fwd;
  if (gid < n_bwd) {
    // transposed images: row n = input feature, k = output feature (or head row)
    int st = gid / per_stage, rem = gid % per_stage, n = rem >> 2, j = rem & 3;
    int layer, kp;
    bwd_stage_info(st, layer, kp);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; e++) {
      int k = kp * 32 + j * 8 + e;
      float x = 0.f;
      if (layer == 9) {
        if (k < 16) {
          const float* w; const float* b; int r;
          head_src(k, w, b, r);
          if (w) x = w[r * 256 + n];
        }
      } else if (layer == 5) {
        x = p.linear_w[5][(size_t)k * (pe_dim + 256) + pe_dim + n];
      } else {
        x = p.linear_w[layer][(size_t)k * 256 + n];
      }
      v[e] = x;
    }
    put(packed + kBwdOff + (size_t)st * kStageBytes + n * 128, n, j, v);
    return;
  }
  gid -= n_bwd;
  if (gid < kBiasFloats) {
    float v = 0.f;
    if (gid < 8 * 256) v = p.linear_b[gid >> 8][gid & 255];
    else {
      const float* w; const float* b; int r;
      head_src(gid - 8 * 256, w, b, r);
      if (b) v = b[r];
    }
    ((float*)(packed + kBiasOff))[gid] = v;
  }
}

// Backward gain exponents (mlp_common.cuh "Backward range management"): block 0..7 = trunk layer l (the h-fed
// columns of its weight), block 8 = the evaluated heads.  e = -round(log2(rms gain of one chain step)).
__global__ void __launch_bounds__(256) pack_gains_kernel(d3g_mlp_config cfg, d3g_mlp_params p, uint8_t* __restrict__ packed) {
  __shared__ float red[8];
  const int l = blockIdx.x, tid = threadIdx.x;
  const int pe_dim = cfg.is_blender ? 93 : 84;
  float acc = 0.f;
  if (l < 8) {
    const float* W = p.linear_w[l];
    const int ld = l == 0 ? pe_dim : (l == 5 ? pe_dim + 256 : 256), col0 = l == 5 ? pe_dim : 0;
    if (l > 0)
      for (int i = tid; i < 256 * 256; i += 256) { const float w = W[(size_t)(i >> 8) * ld + col0 + (i & 255)]; acc += w * w; }
  } else {
    const float* rows[4] = {cfg.is_6dof ? p.branch_w_w : p.warp_w, cfg.is_6dof ? p.branch_v_w : nullptr,
                            (cfg.heads & 2) ? p.rot_w : nullptr, (cfg.heads & 4) ? p.scale_w : nullptr};
    const int nr[4] = {3, 3, 4, 3};
    for (int h = 0; h < 4; h++)
      if (rows[h])
        for (int i = tid; i < nr[h] * 256; i += 256) { const float w = rows[h][i]; acc += w * w; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if ((tid & 31) == 0) red[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) {
    float f2 = 0.f;
    for (int i = 0; i < 8; i++) f2 += red[i];
    // one chain step multiplies a random-direction gradient by ||W||_F / 16 (256 outputs), times sqrt(1/2) for the
    // ReLU mask in the trunk
    const float gain = sqrtf(f2 / 256.0f * (l < 8 ? 0.5f : 1.0f));
    int e = 0;
    if (gain > 0.f && gain < 3.0e38f) e = -(int)rintf(log2f(gain));
    e = e < -12 ? -12 : (e > 12 ? 12 : e);
    ((int*)(packed + kGainOff))[l] = e;
  }
}

int check_cfg(const d3g_mlp_config* cfg) {
  if (!cfg) return set_err(D3G_ERR_INVALID, "cfg is NULL");
  if (cfg->D != 8 || cfg->W != 256 || cfg->multires != 10)
    return set_err(D3G_ERR_UNSUPPORTED, "only D=8, W=256, multires=10 is implemented (got D=%d W=%d multires=%d)",
                   cfg->D, cfg->W, cfg->multires);
  if (!(cfg->heads & 1)) return set_err(D3G_ERR_INVALID, "heads bit0 (warp) must be set");
  return 0;
}

static int g_sm_reserve = 0;     // SMs the persistent MLP kernels leave free (d3g_mlp_sm_reserve)
int mlp_grid_limit() {
  const int n = num_sms() - g_sm_reserve;
  return n < 1 ? 1 : n;
}
static int g_num_sms[64] = {};   // per device
int num_sms() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (!g_num_sms[dev]) {
    cudaDeviceGetAttribute(&g_num_sms[dev], cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms[dev] <= 0) g_num_sms[dev] = 148;
  }
  return g_num_sms[dev];
}

}  // namespace d3g

using namespace d3g;

extern "C" {

int d3g_mlp_packed_bytes(const d3g_mlp_config* cfg, size_t* bytes) {
  if (int e = check_cfg(cfg)) return e;
  if (!bytes) return set_err(D3G_ERR_INVALID, "bytes is NULL");
  *bytes = kPackedBytes;
  return 0;
}

int d3g_mlp_workspace_bytes(const d3g_mlp_config* cfg, int64_t rows, int32_t save, size_t* bytes) {
  if (int e = check_cfg(cfg)) return e;
  if (!bytes || rows < 0) return set_err(D3G_ERR_INVALID, "bad arguments");
  *bytes = mlp_ws_view(nullptr, rows, save != 0).total;
  return 0;
}

int d3g_mlp_workspace_bytes_times(const d3g_mlp_config* cfg, int64_t n_points, int32_t n_times, int32_t save,
                                  size_t* bytes) {
  if (int e = check_cfg(cfg)) return e;
  if (!bytes || n_points < 0 || n_times < 1) return set_err(D3G_ERR_INVALID, "bad arguments");
  *bytes = mlp_ws_view(nullptr, n_points, save != 0, n_times).total;
  return 0;
}

int d3g_mlp_pack_weights(const d3g_mlp_config* cfg, const d3g_mlp_params* p, void* packed, void* stream_) {
  if (int e = check_cfg(cfg)) return e;
  if (!p || !packed) return set_err(D3G_ERR_INVALID, "NULL argument");
  for (int l = 0; l < 8; l++)
    if (!p->linear_w[l] || !p->linear_b[l]) return set_err(D3G_ERR_INVALID, "linear.%d parameters missing", l);
  if (cfg->is_6dof ? (!p->branch_w_w || !p->branch_v_w) : !p->warp_w)
    return set_err(D3G_ERR_INVALID, "warp head parameters missing");
  cudaStream_t stream = (cudaStream_t)stream_;
  int total = (kFwdStages + kBwdStages) * 256 * 4 + kBiasFloats;
  { ProfScope _p(PROF_MLP_PACK, stream);
    pack_weights_kernel<<<(total + 255) / 256, 256, 0, stream>>>(*cfg, *p, (uint8_t*)packed);
    D3G_LAUNCH_CHECK(0, stream);
    pack_gains_kernel<<<9, 256, 0, stream>>>(*cfg, *p, (uint8_t*)packed);
  D3G_LAUNCH_CHECK(0, stream); }
  return 0;
}

// One implementation behind both entry points: `blocks` time blocks of `n` points each (plain call: blocks = 1
// and t / t_stride describe per-row values; times call: t holds one value per block).
static int mlp_forward_impl(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p, int64_t n,
                            int blocks, const float* xyz, const float* t, int64_t t_stride, void* workspace,
                            int32_t save, float* d_xyz, float* d_rot, float* d_scale, cudaStream_t stream) {
  if (int e = check_cfg(cfg)) return e;
  if (!packed || !p || !workspace) return set_err(D3G_ERR_INVALID, "NULL argument");
  if (n < 0 || blocks < 1 || (n > 0 && (!xyz || !t || !d_xyz))) return set_err(D3G_ERR_INVALID, "NULL tensor");
  if (cfg->is_blender && (!p->timenet0_w || !p->timenet0_b || !p->timenet2_w || !p->timenet2_b))
    return set_err(D3G_ERR_INVALID, "timenet parameters missing");
  const MlpShape sh = mlp_shape(n, blocks);
  if ((int64_t)sh.tpb * blocks > ((int64_t)1 << 24)) return set_err(D3G_ERR_CAPACITY, "too many rows");
  MlpWs ws = mlp_ws_view(workspace, n, save != 0, blocks);
  D3G_CUDA_TRY(cudaMemsetAsync(ws.status, 0, 8, stream));
  if (n == 0) return 0;
  const bool per_row = blocks == 1 && t_stride != 0 && n > 1;
  const int64_t rows_t = per_row ? n : blocks;
  const int64_t tstride_eff = per_row ? t_stride : (blocks > 1 ? 1 : 0);
  int* nonuniform = ws.status + 1;
  if (per_row) {
    int g = (int)((n + 255) / 256);
    if (g > 4 * num_sms()) g = 4 * num_sms();
    ProfScope _p(PROF_MLP_TFEAT, stream);
    t_nonuniform_kernel<<<g, 256, 0, stream>>>(n, t, t_stride, nonuniform);
    D3G_LAUNCH_CHECK(0, stream);
  }
  { ProfScope _p(PROF_MLP_TFEAT, stream);
    int64_t g = rows_t < 8 * (int64_t)num_sms() ? rows_t : 8 * (int64_t)num_sms();
    tfeat_kernel<<<(unsigned)g, 256, 0, stream>>>(rows_t, t, tstride_eff, cfg->is_blender, p->timenet0_w,
                                                  p->timenet0_b, p->timenet2_w, p->timenet2_b, ws.tfeat,
                                                  save ? ws.thidden : nullptr, per_row ? nonuniform : nullptr);
  D3G_LAUNCH_CHECK(0, stream); }
  MlpFwdArgs a;
  a.wfwd = (const uint8_t*)packed;
  a.bias = (const float*)((const uint8_t*)packed + kBiasOff);
  a.xyz = xyz; a.rows = n; a.tiles = sh.tiles; a.tpb = sh.tpb;
  a.tfeat = ws.tfeat; a.per_row_t = per_row ? 1 : 0; a.nonuniform = nonuniform;
  a.pe_dim = cfg->is_blender ? 93 : 84; a.t_dim = cfg->is_blender ? 30 : 21;
  a.is_6dof = cfg->is_6dof; a.heads = cfg->heads; a.save = save != 0;
  a.pe_img = ws.pe_img; a.h_img = ws.h_img; a.mask = ws.mask; a.heads_out = ws.heads_out;
  a.d_xyz = d_xyz; a.d_rot = (cfg->heads & 2) ? d_rot : nullptr; a.d_scale = (cfg->heads & 4) ? d_scale : nullptr;
  a.abort_flag = ws.status;
  { const char* e = getenv("D3G_FWD_DBG"); a.dbg = e ? atoi(e) : 0; }
  a.tl = ws.status + 16;
  size_t smem = FwdSmem::total + 1024;
  D3G_CUDA_TRY(cudaFuncSetAttribute(mlp_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = a.tiles < mlp_grid_limit() ? a.tiles : mlp_grid_limit();
  { ProfScope _p(PROF_MLP_FWD, stream);
    mlp_fwd_kernel<<<grid, kFwdThreads, smem, stream>>>(a);
  D3G_LAUNCH_CHECK(0, stream); }
  return 0;
}

int d3g_mlp_forward(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p, int64_t rows,
                    const float* xyz, const float* t, int64_t t_stride, void* workspace, int32_t save,
                    float* d_xyz, float* d_rot, float* d_scale, void* stream_) {
  return mlp_forward_impl(cfg, packed, p, rows, 1, xyz, t, t_stride, workspace, save, d_xyz, d_rot, d_scale,
                          (cudaStream_t)stream_);
}

int d3g_mlp_forward_times(const d3g_mlp_config* cfg, const void* packed, const d3g_mlp_params* p, int64_t n_points,
                          int32_t n_times, const float* xyz, const float* t_values, void* workspace, int32_t save,
                          float* d_xyz, float* d_rot, float* d_scale, void* stream_) {
  if (n_times < 1) return set_err(D3G_ERR_INVALID, "n_times must be >= 1");
  return mlp_forward_impl(cfg, packed, p, n_points, n_times, xyz, t_values, 0, workspace, save, d_xyz, d_rot,
                          d_scale, (cudaStream_t)stream_);
}

int d3g_mlp_sm_reserve(int32_t n_sms) {
  if (n_sms < 0 || n_sms > 64) return set_err(D3G_ERR_INVALID, "sm reserve must be in [0, 64]");
  g_sm_reserve = n_sms;
  return 0;
}

int d3g_mlp_read_status(const void* workspace, int32_t* status_host, void* stream_) {
  if (!workspace || !status_host) return set_err(D3G_ERR_INVALID, "NULL argument");
  cudaStream_t stream = (cudaStream_t)stream_;
  D3G_CUDA_TRY(cudaMemcpyAsync(status_host, workspace, 4, cudaMemcpyDeviceToHost, stream));
  D3G_CUDA_TRY(cudaStreamSynchronize(stream));
  return 0;
}

}  // extern "C"
