// Shared host/device helpers for libd3g.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "d3g.h"

namespace d3g {

// ---- error plumbing (no exceptions cross the C ABI) -------------------------------------------
extern thread_local char g_err[512];
int set_err(int code, const char* fmt, ...);

#define D3G_CUDA_TRY(expr)                                                                   \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return d3g::set_err(D3G_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                          __FILE__, __LINE__);                                               \
  } while (0)

#define D3G_LAUNCH_CHECK(dbg, stream)                                                        \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    if (_e == cudaSuccess && (dbg)) _e = cudaStreamSynchronize(stream);                      \
    if (_e != cudaSuccess)                                                                   \
      return d3g::set_err(D3G_ERR_CUDA, "kernel launch failed: %s (%s:%d)",                  \
                          cudaGetErrorString(_e), __FILE__, __LINE__);                       \
  } while (0)

// ---- optional per-kernel timing (bench.py's roofline leg): CUDA events recorded on the launching stream
// around every kernel of the path while d3g_prof_enable(1) is in force.  Off by default (zero cost).
enum ProfId {
  PROF_PREPROCESS_FWD = 0, PROF_SCAN, PROF_DUPLICATE, PROF_PAD_KEYS, PROF_SORT, PROF_RANGES, PROF_RENDER_FWD,
  PROF_RENDER_BWD, PROF_PREPROCESS_BWD, PROF_MLP_TFEAT, PROF_MLP_FWD, PROF_MLP_DGRAD, PROF_MLP_WGRAD,
  PROF_MLP_FINISH, PROF_MLP_PACK, PROF_LOSS_FWD, PROF_LOSS_BWD, PROF_ADAM, PROF_MLP_DPE, PROF_DENSIFY, PROF_TILE_SORT, PROF_TILE_COUNT, PROF_TILE_SCAN, PROF_TILE_SCATTER, PROF_MLP_AMAX, PROF_COUNT
};
extern bool g_prof_on;
void prof_begin(int id, cudaStream_t stream);
void prof_end(int id, cudaStream_t stream);
struct ProfScope {
  int id; cudaStream_t s;
  ProfScope(int id_, cudaStream_t s_) : id(id_), s(s_) { if (g_prof_on) prof_begin(id, s); }
  ~ProfScope() { if (g_prof_on) prof_end(id, s); }
};

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- explicitly rounded float ops: never contracted, identical to the oracle's C expressions ----
__device__ __forceinline__ float fm(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fa(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fs(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ff(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fdv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fsq(float a) { return __fsqrt_rn(a); }
// C-style min/max (a<b?a:b) so NaN behaviour matches the oracle too
__device__ __forceinline__ float cmin(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float cmax(float a, float b) { return a > b ? a : b; }

// Deterministic exp for x <= 0 — bit-identical to det_exp_f32 in oracle/raster_ref_impl.h.
__device__ __forceinline__ float det_exp(float x) {
  if (x < -87.0f) return 0.0f;
  const float magic = 12582912.0f;
  float t = ff(x, 1.4426950408889634f, magic);
  float n = fs(t, magic);
  float r = ff(n, -0.6931471805599453f, x);
  float p = 1.0f / 720.0f;
  p = ff(p, r, 1.0f / 120.0f);
  p = ff(p, r, 1.0f / 24.0f);
  p = ff(p, r, 1.0f / 6.0f);
  p = ff(p, r, 0.5f);
  p = ff(p, r, 1.0f);
  p = ff(p, r, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

}  // namespace d3g
