// Blackwell (sm_100a) primitives used by the MLP kernels: mbarrier, 1-D bulk async copies (TMA engine,
// UBLKCP), tcgen05 MMA / TMEM alloc / ld / st / commit / fences, and the shared-memory operand
// descriptors.  Everything is inline PTX; no CUTLASS.
//
// Operand images.  Two layouts are used, both pre-arranged in HBM exactly as they sit in shared memory,
// so a tile moves with ONE linear bulk copy and needs no tensor map.
//
// (1) K-major "panelised SWIZZLE_128B image" (forward / input-gradient operands): the matrix X[r][c]
//     (fp32, r = M/N index, c = K index) is cut into panels of 32 columns (=128 B); inside a panel row r
//     sits at byte r*128 and its eight 16-byte chunks are XOR-swizzled with (r & 7):
//         byte(r,c) = (c/32)*panel_stride + r*128 + ((((c%32)/4) ^ (r&7)) * 16) + (c%4)*4
//     = canonical UMMA K-major SW128 layout, SBO = 1024 B, layout type 2.
//
// (2) MN-major "32-byte-base 128B-swizzle image" (weight-gradient operands, K = points): measured on
//     B200 with tools/probe_umma.py — for kind::tf32 an MN-major operand reads zeros with layout types 0
//     and 2 and works with layout type 1 (SWIZZLE_128B_BASE32B): rows of 128 B hold 32 consecutive M/N
//     elements of one k; 4 consecutive k rows form a 512-byte atom whose four 32-byte chunks are
//     XOR-swizzled with (k & 3); SBO = stride between 4-k atoms, LBO = stride between 32-wide M/N panels.
//     A tile of 128 points x C features is stored as [point chunk of 32][panel C/32][32 rows][128 B]:
//         byte(p,f) = (p/32)*(C/32)*4096 + (f/32)*4096 + (p%32)*128 + ((((f%32)/8) ^ (p&3)) * 32) + (f%8)*4
//     so SBO = 512, LBO = 4096 and one 32-point chunk of all C features is one contiguous block.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace d3g {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// byte offset of element (r,c) inside a panelised SW128 image
__device__ __host__ __forceinline__ uint32_t img_off(uint32_t r, uint32_t c, uint32_t panel_stride) {
  return (c >> 5) * panel_stride + r * 128u + ((((c & 31u) >> 2) ^ (r & 7u)) << 4) + ((c & 3u) << 2);
}

// byte offset of (point p, feature f) inside an MN-major wgrad image of a 128 x C tile
__device__ __host__ __forceinline__ uint32_t img_off_mn(uint32_t p, uint32_t f, uint32_t C) {
  return (p >> 5) * (C >> 5) * 4096u + (f >> 5) * 4096u + (p & 31u) * 128u + ((((f & 31u) >> 3) ^ (p & 3u)) << 5) +
         ((f & 7u) << 2);
}

// ---- mbarrier --------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug turns into an error flag + early exit instead of a hung GPU.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, volatile int* abort_flag) {
  for (uint32_t spin = 0; spin < (1u << 22); spin++) {
    if (mbar_try_wait(bar, parity)) return true;
    if ((spin & 1023u) == 1023u && *abort_flag) return false;
  }
  *abort_flag = 1;
  return false;
}

// Non-suspending variant for the single MMA-issuing thread: mbarrier.test_wait returns at once (try_wait may park the
// thread; measured ~180 clk per call even on a completed phase, twice per 32-feature chunk on the issue thread).
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_wait_spin(uint64_t* bar, uint32_t parity, volatile int* abort_flag) {
  for (uint32_t spin = 0; spin < (1u << 24); spin++) {
    if (mbar_test_wait(bar, parity)) return true;
    if ((spin & 4095u) == 4095u && *abort_flag) return false;
  }
  *abort_flag = 1;
  return false;
}

// make generic-proxy writes to shared memory visible to the async proxy (bulk copies, tcgen05.mma)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- 1-D bulk async copies (TMA engine) -------------------------------------------------------
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ---- TMEM allocation (one warp, .sync.aligned) -------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- descriptors -------------------------------------------------------------------------------
// Shared-memory matrix descriptor, sm_100 version bits; layout type 2 = SWIZZLE_128B (K-major images),
// 1 = SWIZZLE_128B_BASE32B (MN-major tf32 images).
__device__ __forceinline__ uint64_t make_smem_desc_t(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                     uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)layout_type << 61;
  return d;
}
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return make_smem_desc_t(saddr, lbo_bytes, sbo_bytes, 2u);
}
// advance the start address of a descriptor by `bytes` (a multiple of 16; shared-memory addresses stay below 2^18, so
// the 14-bit field never carries)
__device__ __forceinline__ uint64_t desc_adv(uint64_t d, uint32_t bytes) { return d + (uint64_t)(bytes >> 4); }
// MN-major wgrad image operand: start of a group of 8 points inside a 32-point chunk
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t saddr) {
  return make_smem_desc_t(saddr, 4096u, 512u, 1u);
}
// kind::tf32 instruction descriptor: D=f32, A=B=tf32, dense; majors: 0 = K-major, 1 = MN-major.
__device__ __host__ __forceinline__ uint32_t make_idesc_tf32(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// kind::f16 instruction descriptor: D=f32, A=B=fp16, dense; majors: 0 = K-major, 1 = MN-major.
__device__ __host__ __forceinline__ uint32_t make_idesc_f16(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
  return (1u << 4) | (0u << 7) | (0u << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// byte offset of element (r,c) of a 16-bit matrix inside a panelised SW128 image: panels of 64 columns (=128 B),
// row r at byte r*128, eight 16-byte chunks XOR-swizzled with (r & 7).  With r = M/N index and c = K index this is
// the canonical K-major SWIZZLE_128B operand (SBO = 1024); the SAME bytes, read with r = K index and c = M/N index,
// are the canonical MN-major SWIZZLE_128B operand (LBO = panel stride, SBO = 1024) — so an activation tile stored
// [point][feature] feeds the layer chain (K = features) and the weight gradient (K = points) without re-arrangement.
__device__ __host__ __forceinline__ uint32_t img_off16(uint32_t r, uint32_t c, uint32_t panel_stride) {
  return (c >> 6) * panel_stride + r * 128u + ((((c & 63u) >> 3) ^ (r & 7u)) << 4) + ((c & 7u) << 1);
}

// ---- MMA issue (one thread) --------------------------------------------------------------------
__device__ __forceinline__ void umma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// ---- TMEM <-> registers: each thread of the warp touches its own lane, 32 consecutive columns ----
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,"
      "%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]),
        "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]),
        "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// round-to-nearest fp32 -> tf32 (the tensor core otherwise truncates the low 13 mantissa bits)
__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

// explicit shared-window accesses: the dynamic-smem pointers reach the epilogues as generic pointers, and the
// generic LD/ST forms cost an extra R2UR pair each
__device__ __forceinline__ void sts128(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float4 lds128f(uint32_t saddr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(saddr));
  return r;
}

// Same rounding on the bit pattern, for finite inputs: add half an ulp of the 10-bit mantissa, drop the low 13 bits
// (magnitude rounding, ties away from zero = cvt.rna).  One integer add + one logic op instead of the guarded
// sequence the compiler emits for cvt.rna.tf32 — the epilogues are ALU-pipe bound.  Inf/NaN stay non-finite.
__device__ __forceinline__ uint32_t tf32_rna_bits(uint32_t bits) { return (bits + 0x1000u) & 0xFFFFE000u; }

}  // namespace d3g
