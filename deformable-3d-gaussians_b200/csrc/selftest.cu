// tcgen05 self-test: one CTA computes D[128,N] = A[128,K] * B[N,K]^T through exactly the operand
// conventions the MLP kernels use (panelised SWIZZLE_128B images, kind::tf32, accumulators in TMEM).
// Called from tests/ through d3g_selftest_umma; it exists so that a wrong descriptor bit shows up as
// a small, localised GEMM failure instead of a wrong MLP.
#include "common.cuh"
#include "d3g_selftest.h"
#include "umma.cuh"

namespace d3g {

// mode 0: A,B K-major in smem (A given [128,K], B given [N,K])
// mode 1: A in TMEM (written with tcgen05.st), B K-major in smem
// mode 2: A,B MN-major in smem (A given [K,128], B given [K,N]), 32-byte-base swizzle images
// mode 3: A MN-major, B = all-ones K-major with N = 16 (the bias-gradient column-sum trick)
__global__ void __launch_bounds__(128, 1)
selftest_umma_kernel(int mode, int N, int K, const float* __restrict__ A, const float* __restrict__ B,
                     float* __restrict__ D, int* abort_flag) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t kpanels = (K + 31) / 32;
  uint8_t* sA = smem;
  uint32_t a_panel_stride, b_panel_stride, a_bytes;
  if (mode >= 2) {
    // MN-major wgrad images: [chunk of 32 points][panel][32 rows][128 B]; K padded up to chunks of 32
    a_panel_stride = 4096u;
    a_bytes = ((uint32_t)(K + 31) / 32u) * 4u * 4096u;
    b_panel_stride = 4096u;
  } else {
    a_panel_stride = 128u * 128u;
    a_bytes = kpanels * a_panel_stride;
    b_panel_stride = (uint32_t)N * 128u;
  }
  uint8_t* sB = sA + ((a_bytes + 1023u) & ~1023u);

  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  // stage operands
  if (mode >= 2) {
    for (int i = tid; i < K * 128; i += 128) {
      int k = i / 128, m = i % 128;
      *(uint32_t*)(sA + img_off_mn(k, m, 128)) = to_tf32(A[(size_t)k * 128 + m]);
    }
    if (mode == 2) {
      for (int i = tid; i < K * N; i += 128) {
        int k = i / N, n = i % N;
        *(uint32_t*)(sB + img_off_mn(k, n, N)) = to_tf32(B[(size_t)k * N + n]);
      }
    } else {
      for (int i = tid; i < 16 * 32; i += 128) ((float*)sB)[i] = 1.0f;  // ones, K-major, 16 rows x 128 B
    }
  } else {
    if (mode == 0)
      for (int i = tid; i < 128 * K; i += 128) {
        int r = i / K, c = i % K;
        *(uint32_t*)(sA + img_off(r, c, a_panel_stride)) = to_tf32(A[(size_t)r * K + c]);
      }
    for (int i = tid; i < N * K; i += 128) {
      int r = i / K, c = i % K;
      *(uint32_t*)(sB + img_off(r, c, b_panel_stride)) = to_tf32(B[(size_t)r * K + c]);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t a_col0 = 256;
  if (mode == 1) {
    // every thread owns one row of A; K is a multiple of 32 in this mode
    for (int c0 = 0; c0 < K; c0 += 32) {
      uint32_t v[32];
#pragma unroll
      for (int j = 0; j < 32; j++) v[j] = to_tf32(A[(size_t)tid * K + c0 + j]);
      tmem_st32(lane_addr + a_col0 + c0, v);
    }
    tmem_wait_st();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (tid == 0) {
    const uint32_t a_mn = mode >= 2, b_mn = mode == 2;
    const uint32_t idesc = make_idesc_tf32(128, N, a_mn, b_mn);
    for (int k8 = 0; k8 < K / 8; k8++) {
      uint64_t bdesc, adesc = 0;
      if (mode == 2) {
        adesc = make_smem_desc_mn(smem_u32(sA) + (k8 / 4) * 4 * 4096 + (k8 % 4) * 1024);
        bdesc = make_smem_desc_mn(smem_u32(sB) + (k8 / 4) * (N / 32) * 4096 + (k8 % 4) * 1024);
      } else if (mode == 3) {
        adesc = make_smem_desc_mn(smem_u32(sA) + (k8 / 4) * 4 * 4096 + (k8 % 4) * 1024);
        bdesc = make_smem_desc(smem_u32(sB), 16, 1024);
      } else {
        adesc = make_smem_desc(smem_u32(sA) + (k8 / 4) * a_panel_stride + (k8 % 4) * 32, 16, 1024);
        bdesc = make_smem_desc(smem_u32(sB) + (k8 / 4) * b_panel_stride + (k8 % 4) * 32, 16, 1024);
      }
      if (mode == 1) umma_tf32_ts(tmem_base, tmem_base + a_col0 + k8 * 8, bdesc, idesc, k8 > 0);
      else umma_tf32_ss(tmem_base, adesc, bdesc, idesc, k8 > 0);
    }
    umma_commit(&bar);
  }
  bool ok = mbar_wait(&bar, 0, abort_flag);
  tc_fence_after();
  if (ok) {
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(lane_addr + c0, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; j++)
        if (c0 + j < N) D[(size_t)tid * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

}  // namespace d3g

using namespace d3g;

extern "C" int d3g_selftest_umma(int32_t mode, int32_t N, int32_t K, const float* A, const float* B, float* D,
                                 void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (mode < 0 || mode > 3) return set_err(D3G_ERR_INVALID, "mode must be 0..3");
  if (mode == 3 && N != 16) return set_err(D3G_ERR_INVALID, "mode 3 needs N=16");
  if (((N < 32 || N % 32) && mode != 3) || N > 256 || K < 8 || K > 64 || K % 8 || (mode == 1 && K % 32))
    return set_err(D3G_ERR_INVALID, "unsupported self-test shape N=%d K=%d", N, K);
  int* flag = nullptr;
  D3G_CUDA_TRY(cudaMalloc(&flag, sizeof(int)));
  D3G_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), stream));
  size_t smem = 1024 + 4 * 128 * 128 + 8 * 64 * 128 * 4 + 2048;
  smem = 1024 + 65536 + 65536 + 2048;
  D3G_CUDA_TRY(cudaFuncSetAttribute(selftest_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  selftest_umma_kernel<<<1, 128, smem, stream>>>(mode, N, K, A, B, D, flag);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  int h = 0;
  if (e == cudaSuccess) e = cudaMemcpy(&h, flag, sizeof(int), cudaMemcpyDeviceToHost);
  cudaFree(flag);
  if (e != cudaSuccess) return set_err(D3G_ERR_CUDA, "self-test failed: %s", cudaGetErrorString(e));
  if (h) return set_err(D3G_ERR_CUDA, "self-test: mbarrier wait timed out (protocol error)");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// kind::f16 self-test (fp16 operands, fp32 accumulation) through the 16-bit operand conventions of umma.cuh:
//   mode 10: A,B K-major SW128 images in shared memory (A given [128,K], B given [N,K])
//   mode 11: A in tensor memory, two consecutive k packed per 32-bit column (low half = even k), B K-major
//   mode 12: A,B MN-major from [K,128] / [K,N] matrices stored with the SAME image formula (rows = k)
// Inputs are fp32 arrays holding fp16-representable values.
// ------------------------------------------------------------------------------------------------
#include <cuda_fp16.h>
namespace d3g {
__global__ void __launch_bounds__(128, 1)
selftest_f16_kernel(int mode, int N, int K, const float* __restrict__ A, const float* __restrict__ B,
                    float* __restrict__ D, int* abort_flag) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  uint8_t* sA = smem;
  uint8_t* sB = smem + 32768;
  const uint32_t a_panel = mode == 12 ? (uint32_t)K * 128u : 128u * 128u;
  const uint32_t b_panel = mode == 12 ? (uint32_t)K * 128u : (uint32_t)N * 128u;
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (mode == 12) {
    for (int i = tid; i < K * 128; i += 128) {
      int k = i / 128, m = i % 128;
      *(__half*)(sA + img_off16(k, m, a_panel)) = __float2half_rn(A[(size_t)k * 128 + m]);
    }
    for (int i = tid; i < K * N; i += 128) {
      int k = i / N, n = i % N;
      *(__half*)(sB + img_off16(k, n, b_panel)) = __float2half_rn(B[(size_t)k * N + n]);
    }
  } else {
    if (mode == 10)
      for (int i = tid; i < 128 * K; i += 128) {
        int r = i / K, c = i % K;
        *(__half*)(sA + img_off16(r, c, a_panel)) = __float2half_rn(A[(size_t)r * K + c]);
      }
    for (int i = tid; i < N * K; i += 128) {
      int r = i / K, c = i % K;
      *(__half*)(sB + img_off16(r, c, b_panel)) = __float2half_rn(B[(size_t)r * K + c]);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t a_col0 = 256;
  if (mode == 11) {
    for (int c0 = 0; c0 < K; c0 += 32) {   // 32 k = 16 packed columns
      uint32_t v[16];
#pragma unroll
      for (int j = 0; j < 16; j++) {
        const __half2 h = __floats2half2_rn(A[(size_t)tid * K + c0 + 2 * j], A[(size_t)tid * K + c0 + 2 * j + 1]);
        v[j] = *(const uint32_t*)&h;
      }
      tmem_st16(lane_addr + a_col0 + c0 / 2, v);
    }
    tmem_wait_st();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (tid == 0) {
    const uint32_t mn = mode == 12;
    const uint32_t idesc = make_idesc_f16(128, N, mn, mn);
    for (int k16 = 0; k16 < K / 16; k16++) {
      if (mode == 12) {
        const uint64_t ad = make_smem_desc_t(smem_u32(sA) + k16 * 2048, a_panel, 1024, 2u);
        const uint64_t bd = make_smem_desc_t(smem_u32(sB) + k16 * 2048, b_panel, 1024, 2u);
        umma_f16_ss(tmem_base, ad, bd, idesc, k16 > 0);
      } else {
        const uint64_t bd = make_smem_desc(smem_u32(sB) + (k16 / 4) * b_panel + (k16 % 4) * 32, 16, 1024);
        if (mode == 11) umma_f16_ts(tmem_base, tmem_base + a_col0 + k16 * 8, bd, idesc, k16 > 0);
        else umma_f16_ss(tmem_base, make_smem_desc(smem_u32(sA) + (k16 / 4) * a_panel + (k16 % 4) * 32, 16, 1024), bd, idesc, k16 > 0);
      }
    }
    umma_commit(&bar);
  }
  bool ok = mbar_wait(&bar, 0, abort_flag);
  tc_fence_after();
  if (ok) {
    for (int c0 = 0; c0 < N; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(lane_addr + c0, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 16; j++)
        if (c0 + j < N) D[(size_t)tid * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}
}  // namespace d3g

extern "C" int d3g_selftest_umma_f16(int32_t mode, int32_t N, int32_t K, const float* A, const float* B, float* D,
                                     void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (mode < 10 || mode > 12) return set_err(D3G_ERR_INVALID, "mode must be 10..12");
  if (N < 16 || N % 16 || N > 256 || K < 16 || K > 128 || K % 16 || (mode == 11 && K % 32) || (mode == 12 && N % 64))
    return set_err(D3G_ERR_INVALID, "unsupported f16 self-test shape N=%d K=%d", N, K);
  int* flag = nullptr;
  D3G_CUDA_TRY(cudaMalloc(&flag, sizeof(int)));
  D3G_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), stream));
  size_t smem = 1024 + 32768 + 65536 + 1024;
  D3G_CUDA_TRY(cudaFuncSetAttribute(selftest_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  selftest_f16_kernel<<<1, 128, smem, stream>>>(mode, N, K, A, B, D, flag);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  int h = 0;
  if (e == cudaSuccess) e = cudaMemcpy(&h, flag, sizeof(int), cudaMemcpyDeviceToHost);
  cudaFree(flag);
  if (e != cudaSuccess) return set_err(D3G_ERR_CUDA, "f16 self-test failed: %s", cudaGetErrorString(e));
  if (h) return set_err(D3G_ERR_CUDA, "f16 self-test: mbarrier wait timed out (protocol error)");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Raw probe: the host supplies byte images for A and B, one (A,B) descriptor pair per MMA (start
// address fields relative to the image base) and the instruction descriptor.  Used by
// tests/test_gpu_umma.py to pin the shared-memory operand conventions empirically.
// ------------------------------------------------------------------------------------------------
namespace d3g {
__global__ void __launch_bounds__(128, 1)
selftest_raw_kernel(const uint8_t* __restrict__ a_img, uint32_t a_bytes, const uint8_t* __restrict__ b_img,
                    uint32_t b_bytes, int n_mma, const uint64_t* __restrict__ adesc,
                    const uint64_t* __restrict__ bdesc, uint32_t idesc, int N, float* __restrict__ D,
                    int* abort_flag, int reps) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  uint8_t* sA = smem;
  uint8_t* sB = sA + ((a_bytes + 1023u) & ~1023u);
  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  for (uint32_t i = tid; i < a_bytes / 4; i += 128) ((uint32_t*)sA)[i] = ((const uint32_t*)a_img)[i];
  for (uint32_t i = tid; i < b_bytes / 4; i += 128) ((uint32_t*)sB)[i] = ((const uint32_t*)b_img)[i];
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (tid == 0) {
    for (int r = 0; r < reps; r++)
      for (int i = 0; i < n_mma; i++) {
        uint64_t ad = adesc[i] + (uint64_t)((smem_u32(sA) >> 4) & 0x3FFFu);
        uint64_t bd = bdesc[i] + (uint64_t)((smem_u32(sB) >> 4) & 0x3FFFu);
        umma_tf32_ss(tmem_base, ad, bd, idesc, (i | r) > 0);
      }
    umma_commit(&bar);
  }
  bool ok = false;
  for (int w = 0; w < 64 && !ok && !*abort_flag; w++) { *abort_flag = 0; ok = mbar_wait(&bar, 0, abort_flag); }
  if (ok) *abort_flag = 0;
  tc_fence_after();
  if (ok) {
    const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(lane_addr + c0, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; j++)
        if (c0 + j < N) D[(size_t)tid * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 256);
}
}  // namespace d3g

extern "C" int d3g_selftest_umma_raw(const void* a_img, uint32_t a_bytes, const void* b_img, uint32_t b_bytes,
                                     int32_t n_mma, const uint64_t* adesc, const uint64_t* bdesc, uint32_t idesc,
                                     int32_t N, float* D, int32_t reps, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a_bytes % 4 || b_bytes % 4 || a_bytes > 98304 || b_bytes > 98304 || N < 8 || N > 256 || n_mma < 1)
    return set_err(D3G_ERR_INVALID, "unsupported raw self-test arguments");
  int* flag = nullptr;
  D3G_CUDA_TRY(cudaMalloc(&flag, sizeof(int)));
  D3G_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), stream));
  size_t smem = 1024 + 98304 + 98304 + 1024;
  D3G_CUDA_TRY(cudaFuncSetAttribute(selftest_raw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  selftest_raw_kernel<<<1, 128, smem, stream>>>((const uint8_t*)a_img, a_bytes, (const uint8_t*)b_img, b_bytes, n_mma,
                                                adesc, bdesc, idesc, N, D, flag, reps < 1 ? 1 : reps);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  int h = 0;
  if (e == cudaSuccess) e = cudaMemcpy(&h, flag, sizeof(int), cudaMemcpyDeviceToHost);
  cudaFree(flag);
  if (e != cudaSuccess) return set_err(D3G_ERR_CUDA, "raw self-test failed: %s", cudaGetErrorString(e));
  if (h) return set_err(D3G_ERR_CUDA, "raw self-test: mbarrier wait timed out");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Throughput probe for the MLP's inner step: `reps` x 4 MMAs (M=128, N=256, K=8, kind::tf32) issued by one
// thread, timed with clock64 from first issue to commit arrival.
//   mode 0: A from shared memory (SS)      mode 1: A from tensor memory (TS, as the trunk layers use it)
//   mode 2: TS while warps 1..3 + 4..7 run the epilogue's tcgen05.ld/st pattern on the accumulator half
//   mode 3: SS with the same interference
//   mode 4: TS while eight warps spin on an mbarrier with all 32 lanes (what idle epilogue warps used to do)
//   +8: tcgen05.commit after every 4 MMAs (as the weight ring does)   +16: non-zero operand data
// Reports clocks per MMA per CTA (grid = any number of CTAs, one per SM).
// ------------------------------------------------------------------------------------------------
namespace d3g {
__global__ void __launch_bounds__(288, 1)
umma_clk_kernel(int mode_flags, int reps, long long* __restrict__ out, int* abort_flag) {
  const int mode = mode_flags & 7;
  const bool commit_each = mode_flags & 8, nonzero = mode_flags & 16;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int stop;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (warp == 8) tmem_alloc(&tmem_base_s, 512);
  if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); stop = 0; }
  __shared__ uint64_t bar2;
  if (tid == 0) mbar_init(&bar2, 1);
  for (uint32_t i = tid; i < (65536u + 32768u) / 4; i += 288)
    ((uint32_t*)smem)[i] = nonzero ? to_tf32(0.001f * (float)((i * 2654435761u) >> 20) - 2.0f) : 0u;
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (nonzero && warp < 4) {
    uint32_t v[32];
#pragma unroll
    for (int j = 0; j < 32; j++) v[j] = to_tf32(0.37f + 0.01f * j + 0.001f * lane);
    for (int c = 0; c < 16; c++) tmem_st32(tmem_base + ((uint32_t)(warp * 32) << 16) + c * 32, v);
    tmem_wait_st();
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  if (warp == 8) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, 256, 0, 0);
      const uint32_t sa = smem_u32(smem), sb = smem_u32(smem + 65536);
      long long t0 = clock64();
      for (int r = 0; r < reps; r++) {
#pragma unroll
        for (int ks = 0; ks < 4; ks++) {
          if (mode & 1)
            umma_tf32_ts(tmem_base, tmem_base + 256 + (r & 7) * 32 + ks * 8, make_smem_desc(sb + ks * 32, 16, 1024), idesc, 1u);
          else
            umma_tf32_ss(tmem_base, make_smem_desc(sa + ks * 32, 16, 1024), make_smem_desc(sb + ks * 32, 16, 1024), idesc, 1u);
        }
        if (commit_each) umma_commit(&bar2);
      }
      umma_commit(&bar);
      mbar_wait(&bar, 0, abort_flag);
      long long t1 = clock64();
      out[blockIdx.x] = t1 - t0;
      stop = 1;
    }
  } else if (mode == 4) {
    // eight warps spinning on an mbarrier that never completes (what idle epilogue warps do)
    __shared__ uint64_t never;
    if (tid == 0) mbar_init(&never, 1);
    __syncwarp();
    while (!stop) mbar_try_wait(&never, 0);
  } else if (mode >= 2) {
    // epilogue-like traffic: 2 warps per lane quarter, ld 32 columns -> a few ALU ops -> st back, on columns 256..511
    const uint32_t lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 256;
    int c = warp >> 2;
    while (!stop) {
      uint32_t v[32];
      tmem_ld32(lane_addr + c * 32, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; j++) v[j] = to_tf32(fmaxf(__uint_as_float(v[j]) + 1.0f, 0.f));
      tmem_st32(lane_addr + c * 32, v);
      tmem_wait_st();
      c = (c + 2) & 7;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem_base, 512);
}
}  // namespace d3g

extern "C" int d3g_selftest_umma_clk(int32_t mode, int32_t reps, int32_t ctas, int64_t* clk_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (mode < 0 || mode > 31 || (mode & 7) > 4 || reps < 1 || ctas < 1 || ctas > 1024 || !clk_out)
    return set_err(D3G_ERR_INVALID, "bad umma_clk arguments");
  int* flag = nullptr;
  D3G_CUDA_TRY(cudaMalloc(&flag, sizeof(int)));
  D3G_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), stream));
  size_t smem = 1024 + 65536 + 32768;
  D3G_CUDA_TRY(cudaFuncSetAttribute(umma_clk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  umma_clk_kernel<<<ctas, 288, smem, stream>>>(mode, reps, (long long*)clk_out, flag);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  cudaFree(flag);
  if (e != cudaSuccess) return set_err(D3G_ERR_CUDA, "umma_clk failed: %s", cudaGetErrorString(e));
  return 0;
}
