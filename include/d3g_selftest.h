/*
 * d3g_selftest.h — tcgen05 probes of libd3g.so.  TEST / TOOLING ONLY: these entry points exist so that a wrong
 * descriptor bit shows up as a small, localised GEMM failure (tests/test_gpu_umma.py) and so that instruction
 * throughput can be measured (tools/bench_umma.py).  They are not part of the product ABI (include/d3g.h) and
 * nothing on the hot path calls them.
 */
#ifndef D3G_SELFTEST_H_
#define D3G_SELFTEST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* tcgen05 self-test used by tests/: D[128,N] = A[128,K] * B[N,K]^T through the exact smem/TMEM
 * operand conventions the MLP kernels use.  mode 0: A,B K-major from smem; 1: A from TMEM;
 * 2: A,B MN-major from smem (A given as [K,128], B as [K,N]); 3: A MN-major, B = ones (N = 16), i.e.
 * D[m,:] = sum_k A[k,m] (the bias-gradient trick). */
int d3g_selftest_umma(int32_t mode, int32_t N, int32_t K, const float* A, const float* B, float* D,
                      void* stream);

/* Raw tcgen05 probe (tests / tools only): host-supplied operand images, one shared-memory descriptor
 * pair per MMA (start-address field relative to the image), host-supplied instruction descriptor.
 * D is [128,N].  The MMA list is issued `reps` times (reps > 1 is used to time instruction throughput). */
int d3g_selftest_umma_raw(const void* a_img, uint32_t a_bytes, const void* b_img, uint32_t b_bytes,
                          int32_t n_mma, const uint64_t* adesc, const uint64_t* bdesc, uint32_t idesc,
                          int32_t N, float* D, int32_t reps, void* stream);

/* Throughput probe: `reps` x 4 MMAs (M=128, N=256, K=8, kind::tf32) per CTA on `ctas` CTAs; mode 0 = A from
 * shared memory, 1 = A from tensor memory, 2/3 = the same with epilogue-like tcgen05.ld/st traffic from eight
 * other warps, 4 = A from tensor memory while eight warps spin on an mbarrier.  clk_out (device, int64[ctas]) receives the clocks from first issue to commit arrival. */
int d3g_selftest_umma_clk(int32_t mode, int32_t reps, int32_t ctas, int64_t* clk_out, void* stream);

/* kind::f16 counterpart (fp16 operands, fp32 accumulation) through the 16-bit operand conventions of the MLP
 * kernels.  mode 10: A,B K-major SWIZZLE_128B images from shared memory (A [128,K], B [N,K]); 11: A from tensor
 * memory, two consecutive k packed per column; 12: A,B MN-major (A given as [K,128], B as [K,N]) read from the very
 * same image layout with rows = k. */
int d3g_selftest_umma_f16(int32_t mode, int32_t N, int32_t K, const float* A, const float* B, float* D,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* D3G_SELFTEST_H_ */
