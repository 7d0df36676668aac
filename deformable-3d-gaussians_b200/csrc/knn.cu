// `simple_knn._C.distCUDA2`: mean squared distance of every point to its three nearest
// neighbours, the initial-scale heuristic of `GaussianModel.create_from_pcd` (scene/gaussian_model.py:105).
// Reference: submodules/simple-knn/simple_knn.cu:185-221 (`SimpleKNN::knn`: Morton sort, 1024-point boxes, box-pruned
// exact 3-NN) behind submodules/simple-knn/spatial.cu:15-26.  The reference's search is EXACT (pruning only skips
// boxes that cannot hold a closer point), so the result is the multiset of the three smallest squared distances to
// the OTHER indices (a duplicate point at another index counts with distance 0), summed smallest-first and divided
// by 3.0f — independent of the traversal order.  This file computes the same value with no sort, no host
// round-trips (the reference does two blocking cudaMemcpy for the bounding box) and no temporary allocations:
// every thread owns one query and streams all P points through shared memory in 16-byte broadcast reads; the P
// points (12 B each) stay in L2.  Bit-exactness with the reference rests on the distance being the same expression
// (d = other - query; d.x*d.x + d.y*d.y + d.z*d.z, which nvcc contracts to fma(dz,dz, fma(dx,dx, dy*dy)) — SASS of
// both kernels checked with cuobjdump; written with explicit intrinsics here) — tests/test_gpu_knn.py
// compares with the reference compiled into oracle/_ref/.
// Cost: P^2 distance evaluations (~8 issue slots each): 100k points ~ 2-3 ms, init-only.
#include <cfloat>
#include "common.cuh"

namespace d3g {

constexpr int kKnnThreads = 128;
constexpr int kKnnTile = 1024;   // points per shared-memory tile (16 KB as float4)

__device__ __forceinline__ void knn_insert(float dist, float& b0, float& b1, float& b2) {
  // reference updateKBest<3>: strict `knn[j] > dist` bubble (simple_knn.cu:141-155)
  if (b0 > dist) { const float t = b0; b0 = dist; dist = t; }
  if (b1 > dist) { const float t = b1; b1 = dist; dist = t; }
  if (b2 > dist) { b2 = dist; }
}

__global__ void __launch_bounds__(kKnnThreads)
knn_mean_dist2_kernel(int P, const float* __restrict__ pts, float* __restrict__ out) {
  __shared__ float4 tile[kKnnTile];
  const int i = blockIdx.x * kKnnThreads + threadIdx.x;
  const bool live = i < P;
  const float3 q = live ? make_float3(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2])
                        : make_float3(0.f, 0.f, 0.f);
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  for (int base = 0; base < P; base += kKnnTile) {
    const int cnt = min(kKnnTile, P - base);
    __syncthreads();
    // coalesced fill: the tile's 3*cnt floats are contiguous in global memory
    for (int e = threadIdx.x; e < 3 * cnt; e += kKnnThreads)
      ((float*)tile)[(e / 3) * 4 + (e % 3)] = pts[3 * (size_t)base + e];
    __syncthreads();
    if (live) {
      const int self = i - base;   // index of the query inside this tile (or outside [0, cnt))
#pragma unroll 8
      for (int j = 0; j < cnt; j++) {
        const float4 p = tile[j];
        const float3 d = {p.x - q.x, p.y - q.y, p.z - q.z};
        const float dist = ff(d.z, d.z, ff(d.x, d.x, fm(d.y, d.y)));   // nvcc 12.9's contraction of the reference's sum
        if (b2 > dist && j != self) knn_insert(dist, b0, b1, b2);
      }
    }
  }
  if (live) out[i] = (b0 + b1 + b2) / 3.0f;
}

}  // namespace d3g

using namespace d3g;

extern "C" int d3g_knn_mean_dist2(int32_t P, const float* points, float* mean_dist2, void* stream_) {
  if (P < 0 || (P > 0 && (!points || !mean_dist2))) return set_err(D3G_ERR_INVALID, "bad knn arguments");
  if (P == 0) return 0;
  cudaStream_t stream = (cudaStream_t)stream_;
  knn_mean_dist2_kernel<<<(P + kKnnThreads - 1) / kKnnThreads, kKnnThreads, 0, stream>>>(P, points, mean_dist2);
  D3G_LAUNCH_CHECK(0, stream);
  return 0;
}
