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
#include <cub/device/device_scan.cuh>
#include "common.cuh"

namespace d3g {

constexpr int kKnnThreads = 128;
constexpr int kKnnTile = 1024;   // points per shared-memory tile (16 KB as float4)
constexpr int kKnnChunk = 8;     // distances evaluated between two insertion checks (kKnnTile % kKnnChunk == 0)

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
    // pad the tail of the last tile to a multiple of the chunk with +inf points: their distance is +inf, never < b2
    for (int e = cnt + threadIdx.x; e < ((cnt + kKnnChunk - 1) / kKnnChunk) * kKnnChunk; e += kKnnThreads)
      tile[e] = make_float4(INFINITY, INFINITY, INFINITY, 0.f);
    __syncthreads();
    {                              // threads past P run along with q = 0 (full-warp vote below); only the store is guarded
      const int self = i - base;   // index of the query inside this tile (or outside [0, cnt))
      for (int j = 0; j < cnt; j += kKnnChunk) {
        // fast path: kKnnChunk distances and their minimum, six FP32 instructions + one FMNMX per pair; the
        // (rare after the first tiles) insertion is a real, warp-uniform branch instead of ~12 predicated
        // instructions per pair.  The query's own distance 0 only ever triggers the slow path, which skips it.
        float dist[kKnnChunk];
        float m = FLT_MAX;
#pragma unroll
        for (int k = 0; k < kKnnChunk; k++) {
          const float4 p = tile[j + k];
          const float3 d = {p.x - q.x, p.y - q.y, p.z - q.z};
          dist[k] = ff(d.z, d.z, ff(d.x, d.x, fm(d.y, d.y)));   // nvcc 12.9's contraction of the reference's sum
          m = fminf(m, dist[k]);
        }
        if (__any_sync(0xffffffffu, b2 > m)) {
#pragma unroll
          for (int k = 0; k < kKnnChunk; k++)
            if (b2 > dist[k] && j + k != self) knn_insert(dist[k], b0, b1, b2);
        }
      }
    }
  }
  if (live) out[i] = (b0 + b1 + b2) / 3.0f;
}


// ---------------------------------------------------------------------------------------------------------------
// Large clouds (C3 300k, C5 1M points): exact 3-NN over a uniform grid instead of all P^2 pairs.  Everything runs on
// the caller's stream with no host round-trip: bounding box by atomics on order-preserving integer keys -> a
// one-thread plan (cubic cells of ~4 points) -> counting sort of the points by cell (atomic rank + CUB exclusive
// scan + scatter) -> one thread per sorted point visits Chebyshev shells r = 0, 1, 2, ... of cells around its own
// and stops once the third-best distance is below the smallest distance any unvisited cell can hold.  Cells are
// x-fastest, so the x-run of a shell row is ONE contiguous range of the sorted array.  The search is exact (the
// stop test keeps a 1e-3 relative margin, far above the <= 2e-4-cell error of the float cell assignment), so the
// result equals the brute-force kernel's and the reference's bit for bit.
struct KnnPlan { float minx, miny, minz, h, inv_h; int gx, gy, gz; };
constexpr int kKnnMaxDim = 512;

__device__ __forceinline__ uint32_t knn_enc(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float knn_dec(uint32_t e) {
  return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e);
}

__global__ void __launch_bounds__(256) knn_bbox_kernel(int P, const float* __restrict__ pts, uint32_t* bbox) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x)
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float v = pts[3 * (size_t)i + c];
      lo[c] = fminf(lo[c], v); hi[c] = fmaxf(hi[c], v);
    }
  __shared__ float red[2][3][8];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    for (int o = 16; o; o >>= 1) {
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
    if ((threadIdx.x & 31) == 0) { red[0][c][threadIdx.x >> 5] = lo[c]; red[1][c][threadIdx.x >> 5] = hi[c]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {           // six atomics per CTA (one per bound), not six per warp
    const int which = threadIdx.x / 3, c = threadIdx.x % 3;
    float v = red[which][c][0];
    for (int w = 1; w < 8; w++) v = which ? fmaxf(v, red[1][c][w]) : fminf(v, red[0][c][w]);
    if (which) atomicMax(&bbox[3 + c], knn_enc(v)); else atomicMin(&bbox[c], knn_enc(v));
  }
}

// Robust extent of the cloud.  A few far outliers (COLMAP clouds have them) would stretch a grid laid over the raw
// bounding box until the whole dense part falls into a handful of cells and the query degenerates to all-pairs.  Three
// passes of "mean +- 4 sigma of the points inside the current box" shrink the box onto the bulk; points outside are
// clamped into its boundary cells by knn_cell, which keeps the search exact (a clamped point is at least as far away
// as the cell it is filed under, so the shell bound stays a lower bound).  mom: [count, sum xyz, sum of squares xyz].
__global__ void __launch_bounds__(256)
knn_moments_kernel(int P, const float* __restrict__ pts, const float* __restrict__ box, double* mom) {
  double acc[7] = {0, 0, 0, 0, 0, 0, 0};
  const float lo[3] = {box[0], box[1], box[2]}, hi[3] = {box[3], box[4], box[5]};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
    const float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    if (x >= lo[0] && x <= hi[0] && y >= lo[1] && y <= hi[1] && z >= lo[2] && z <= hi[2]) {
      acc[0] += 1.0; acc[1] += x; acc[2] += y; acc[3] += z;
      acc[4] += (double)x * x; acc[5] += (double)y * y; acc[6] += (double)z * z;
    }
  }
#pragma unroll
  for (int k = 0; k < 7; k++) {
    for (int o = 16; o; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
    if ((threadIdx.x & 31) == 0 && acc[k] != 0.0) atomicAdd(&mom[k], acc[k]);
  }
}
// box <- (mean +- 4 sigma) intersected with box; pass 0 starts from the raw bounding box
__global__ void knn_trim_kernel(const uint32_t* bbox, int first, double* mom, float* box) {
  if (first) {
    for (int c = 0; c < 6; c++) box[c] = knn_dec(bbox[c]);
  } else if (mom[0] >= 8.0) {
    for (int c = 0; c < 3; c++) {
      const double mean = mom[1 + c] / mom[0];
      const double var = fmax(mom[4 + c] / mom[0] - mean * mean, 0.0);
      const double sd = sqrt(var);
      const float lo = (float)(mean - 4.0 * sd), hi = (float)(mean + 4.0 * sd);
      if (lo < hi) { box[c] = fmaxf(box[c], lo); box[3 + c] = fminf(box[3 + c], hi); }
    }
  }
  for (int k = 0; k < 7; k++) mom[k] = 0.0;
}

__global__ void knn_plan_kernel(int P, int cell_cap, const float* box, KnnPlan* plan) {
  KnnPlan pl;
  pl.minx = box[0]; pl.miny = box[1]; pl.minz = box[2];
  const float ex = box[3] - pl.minx, ey = box[4] - pl.miny, ez = box[5] - pl.minz;
  const float me = fmaxf(ex, fmaxf(ey, ez));
  if (!(me > 0.f) || !(me < FLT_MAX)) {          // all points coincide (or non-finite input): one cell
    pl.h = 1.f; pl.inv_h = 0.f; pl.gx = pl.gy = pl.gz = 1;
  } else {
    const float floor_e = me / (float)kKnnMaxDim;   // flat / thin clouds: size the cells by area, not by volume
    const float vol = fmaxf(ex, floor_e) * fmaxf(ey, floor_e) * fmaxf(ez, floor_e);
    float h = fmaxf(cbrtf(vol * 4.f / (float)P), floor_e);
    for (;;) {
      pl.gx = min(kKnnMaxDim, (int)(ex / h) + 1); pl.gy = min(kKnnMaxDim, (int)(ey / h) + 1);
      pl.gz = min(kKnnMaxDim, (int)(ez / h) + 1);
      if ((long long)pl.gx * pl.gy * pl.gz <= (long long)cell_cap) break;
      h *= 1.26f;
    }
    pl.h = h; pl.inv_h = 1.f / h;
  }
  *plan = pl;
}

__device__ __forceinline__ int3 knn_cell(const KnnPlan& pl, float x, float y, float z) {
  return make_int3(min(pl.gx - 1, max(0, (int)((x - pl.minx) * pl.inv_h))),
                   min(pl.gy - 1, max(0, (int)((y - pl.miny) * pl.inv_h))),
                   min(pl.gz - 1, max(0, (int)((z - pl.minz) * pl.inv_h))));
}

__global__ void __launch_bounds__(256)
knn_count_kernel(int P, const float* __restrict__ pts, const KnnPlan* __restrict__ plan, uint32_t* counts,
                 uint32_t* cell_of, uint32_t* rank_of) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const KnnPlan pl = *plan;
  const int3 c = knn_cell(pl, pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2]);
  const uint32_t cell = ((uint32_t)c.z * pl.gy + c.y) * pl.gx + c.x;
  cell_of[i] = cell;
  rank_of[i] = atomicAdd(&counts[cell], 1u);
}

__global__ void __launch_bounds__(256)
knn_scatter_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ starts,
                   const uint32_t* __restrict__ cell_of, const uint32_t* __restrict__ rank_of, float4* sorted) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  sorted[starts[cell_of[i]] + rank_of[i]] =
      make_float4(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], __int_as_float(i));
}

constexpr int kKnnOutlierCap = 4096;

__global__ void __launch_bounds__(128)
knn_grid_query_kernel(int P, const KnnPlan* __restrict__ plan, const uint32_t* __restrict__ starts,
                      const float4* __restrict__ sorted, float* __restrict__ out, const float* __restrict__ box,
                      uint32_t* __restrict__ outliers) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= P) return;
  const KnnPlan pl = *plan;
  const float4 me = sorted[s];
  const int self = __float_as_int(me.w);
  // a point outside the trimmed box has its neighbours far away in grid units: its shell walk would visit the whole
  // grid in one thread.  It goes on a (short) list instead and gets a CTA of its own (knn_outlier_kernel).
  if (me.x < box[0] || me.y < box[1] || me.z < box[2] || me.x > box[3] || me.y > box[4] || me.z > box[5]) {
    const uint32_t slot = atomicAdd(&outliers[0], 1u);
    if (slot < (uint32_t)kKnnOutlierCap) { outliers[1 + slot] = (uint32_t)s; return; }
  }
  const int3 c = knn_cell(pl, me.x, me.y, me.z);
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  auto scan = [&](uint32_t a, uint32_t b) {
    for (uint32_t k = a; k < b; k++) {
      const float4 p = sorted[k];
      const float3 d = {p.x - me.x, p.y - me.y, p.z - me.z};
      const float dist = ff(d.z, d.z, ff(d.x, d.x, fm(d.y, d.y)));
      if (b2 > dist && __float_as_int(p.w) != self) knn_insert(dist, b0, b1, b2);
      if (b2 == 0.0f) return;   // three coincident neighbours: nothing can improve (clouds with many duplicates)
    }
  };
  const int rmax = max(pl.gx, max(pl.gy, pl.gz));
  for (int r = 0; r < rmax; r++) {
    for (int dz = -r; dz <= r; dz++) {
      const int z = c.z + dz;
      if (z < 0 || z >= pl.gz) continue;
      for (int dy = -r; dy <= r; dy++) {
        const int y = c.y + dy;
        if (y < 0 || y >= pl.gy) continue;
        const uint32_t row = ((uint32_t)z * pl.gy + y) * pl.gx;
        if (abs(dz) == r || abs(dy) == r) {       // a face row of the shell: the whole x-run is contiguous
          scan(starts[row + max(0, c.x - r)], starts[row + min(pl.gx - 1, c.x + r) + 1]);
        } else {                                  // interior row: only the two end cells belong to shell r
          if (c.x - r >= 0) scan(starts[row + c.x - r], starts[row + c.x - r + 1]);
          if (c.x + r < pl.gx) scan(starts[row + c.x + r], starts[row + c.x + r + 1]);
        }
      }
    }
    // every unvisited point sits in a cell at Chebyshev offset >= r+1, i.e. at distance >= r*h (minus the cell
    // assignment's rounding, <= 2e-4 h): stop once the third-best distance is safely below that
    const float reach = (float)r * pl.h;
    if (b2 < reach * reach * 0.999f || b2 == 0.0f) break;
  }
  out[self] = (b0 + b1 + b2) / 3.0f;
}

// one CTA per listed outlier: all P points scanned by 256 threads, the per-thread three best merged by thread 0
__global__ void __launch_bounds__(256)
knn_outlier_kernel(int P, const float4* __restrict__ sorted, const uint32_t* __restrict__ outliers,
                   float* __restrict__ out) {
  const uint32_t n = min(outliers[0], (uint32_t)kKnnOutlierCap);
  if (blockIdx.x >= n) return;
  const float4 me = sorted[outliers[1 + blockIdx.x]];
  const int self = __float_as_int(me.w);
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  for (int k = threadIdx.x; k < P; k += blockDim.x) {
    const float4 p = sorted[k];
    const float3 d = {p.x - me.x, p.y - me.y, p.z - me.z};
    const float dist = ff(d.z, d.z, ff(d.x, d.x, fm(d.y, d.y)));
    if (b2 > dist && __float_as_int(p.w) != self) knn_insert(dist, b0, b1, b2);
  }
  __shared__ float cand[3 * 256];
  cand[3 * threadIdx.x] = b0; cand[3 * threadIdx.x + 1] = b1; cand[3 * threadIdx.x + 2] = b2;
  __syncthreads();
  if (threadIdx.x == 0) {
    float c0 = FLT_MAX, c1 = FLT_MAX, c2 = FLT_MAX;
    for (int k = 0; k < 3 * 256; k++)
      if (c2 > cand[k]) knn_insert(cand[k], c0, c1, c2);
    out[self] = (c0 + c1 + c2) / 3.0f;
  }
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

// workspace layout of the grid path (all offsets 256-byte aligned)
namespace {
struct KnnLayout { size_t bbox, box, mom, outliers, plan, counts, starts, cell_of, rank_of, sorted, cub, cub_bytes, total; int cell_cap; };
KnnLayout knn_layout(int32_t P) {
  KnnLayout L;
  L.cell_cap = P / 2 + 4096;
  size_t o = 0;
  auto take = [&](size_t bytes) { const size_t at = o; o = d3g::align_up(o + bytes, 256); return at; };
  L.bbox = take(6 * sizeof(uint32_t));
  L.box = take(6 * sizeof(float));
  L.mom = take(7 * sizeof(double));
  L.outliers = take((size_t)(1 + d3g::kKnnOutlierCap) * sizeof(uint32_t));
  L.plan = take(sizeof(KnnPlan));
  L.counts = take((size_t)(L.cell_cap + 1) * sizeof(uint32_t));
  L.starts = take((size_t)(L.cell_cap + 1) * sizeof(uint32_t));
  L.cell_of = take((size_t)P * sizeof(uint32_t));
  L.rank_of = take((size_t)P * sizeof(uint32_t));
  L.sorted = take((size_t)P * sizeof(float4));
  L.cub_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, L.cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, L.cell_cap + 1);
  L.cub = take(L.cub_bytes);
  L.total = o;
  return L;
}
}  // namespace

extern "C" int d3g_knn_workspace_bytes(int32_t P, size_t* bytes) {
  if (P < 0 || !bytes) return set_err(D3G_ERR_INVALID, "bad knn workspace arguments");
  *bytes = knn_layout(P).total;
  return 0;
}

extern "C" int d3g_knn_mean_dist2_grid(int32_t P, const float* points, float* mean_dist2, void* workspace,
                                       size_t workspace_bytes, void* stream_) {
  if (P < 0 || (P > 0 && (!points || !mean_dist2 || !workspace)))
    return set_err(D3G_ERR_INVALID, "bad knn arguments");
  if (P == 0) return 0;
  const KnnLayout L = knn_layout(P);
  if (workspace_bytes < L.total)
    return set_err(D3G_ERR_CAPACITY, "knn workspace too small: %zu < %zu bytes", workspace_bytes, L.total);
  cudaStream_t stream = (cudaStream_t)stream_;
  char* ws = (char*)workspace;
  uint32_t* bbox = (uint32_t*)(ws + L.bbox);
  KnnPlan* plan = (KnnPlan*)(ws + L.plan);
  uint32_t *counts = (uint32_t*)(ws + L.counts), *starts = (uint32_t*)(ws + L.starts);
  uint32_t *cell_of = (uint32_t*)(ws + L.cell_of), *rank_of = (uint32_t*)(ws + L.rank_of);
  float4* sorted = (float4*)(ws + L.sorted);
  cudaMemsetAsync(bbox, 0xff, 3 * sizeof(uint32_t), stream);        // encoded minima start at the largest key
  cudaMemsetAsync(bbox + 3, 0x00, 3 * sizeof(uint32_t), stream);    // encoded maxima at the smallest
  cudaMemsetAsync(counts, 0, (size_t)(L.cell_cap + 1) * sizeof(uint32_t), stream);
  const int nb = (P + 255) / 256;
  knn_bbox_kernel<<<min(nb, 148 * 4), 256, 0, stream>>>(P, points, bbox);
  float* box = (float*)(ws + L.box);
  double* mom = (double*)(ws + L.mom);
  knn_trim_kernel<<<1, 1, 0, stream>>>(bbox, 1, mom, box);
  for (int pass = 0; pass < 3; pass++) {
    knn_moments_kernel<<<min(nb, 148 * 4), 256, 0, stream>>>(P, points, box, mom);
    knn_trim_kernel<<<1, 1, 0, stream>>>(bbox, 0, mom, box);
  }
  knn_plan_kernel<<<1, 1, 0, stream>>>(P, L.cell_cap, box, plan);
  knn_count_kernel<<<nb, 256, 0, stream>>>(P, points, plan, counts, cell_of, rank_of);
  size_t cub_bytes = L.cub_bytes;
  cudaError_t e = cub::DeviceScan::ExclusiveSum(ws + L.cub, cub_bytes, counts, starts, L.cell_cap + 1, stream);
  if (e != cudaSuccess) return set_err(D3G_ERR_CUDA, "knn scan failed: %s", cudaGetErrorString(e));
  knn_scatter_kernel<<<nb, 256, 0, stream>>>(P, points, starts, cell_of, rank_of, sorted);
  uint32_t* outliers = (uint32_t*)(ws + L.outliers);
  cudaMemsetAsync(outliers, 0, sizeof(uint32_t), stream);
  knn_grid_query_kernel<<<(P + 127) / 128, 128, 0, stream>>>(P, plan, starts, sorted, mean_dist2, box, outliers);
  knn_outlier_kernel<<<kKnnOutlierCap, 256, 0, stream>>>(P, sorted, outliers, mean_dist2);
  D3G_LAUNCH_CHECK(0, stream);
  return 0;
}
