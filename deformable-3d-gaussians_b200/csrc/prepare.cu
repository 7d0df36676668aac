// Gaussian parameter activation + deformation, fused: the ~25 elementwise torch launches that
// gaussian_renderer/__init__.py:69-97 and the getters of scene/gaussian_model.py:58-78,127-146 spend per
// frame between DeformModel.step and the rasterizer become one kernel forward and one backward.
//
//   means3D   = xyz + d_xyz                                  (gaussian_renderer/__init__.py:78)
//             | from_homogenous(d_xyz[4x4] . [xyz,1])         (:75-76, utils/rigid_utils.py:86-107)
//   opacity   = sigmoid(_opacity)                            (:79, gaussian_model.py:63)
//   scales    = exp(_scaling) + d_scaling                    (:96, gaussian_model.py:59)
//   rotations = normalize(_rotation) + d_rotation            (:97, gaussian_model.py:68; F.normalize eps 1e-12)
//
// Pure streaming work: one thread per Gaussian, inputs read once, outputs written once.
#include <math.h>

#include "common.cuh"

namespace d3g {

struct PrepArgs {
  int N;
  int is_6dof;  // d_xyz is [N,4,4] row-major
  const float* xyz; const float* d_xyz;
  const float* scaling; const float* d_scaling;
  const float* rotation; const float* d_rotation;
  const float* opacity;
};

__global__ void __launch_bounds__(256)
prepare_fwd_kernel(PrepArgs a, float* __restrict__ means3D, float* __restrict__ opac,
                   float* __restrict__ scales, float* __restrict__ rots) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.N) return;
  const size_t si = (size_t)i;
  const float x = a.xyz[3 * si], y = a.xyz[3 * si + 1], z = a.xyz[3 * si + 2];
  if (!a.is_6dof) {
    means3D[3 * si] = x + a.d_xyz[3 * si];
    means3D[3 * si + 1] = y + a.d_xyz[3 * si + 1];
    means3D[3 * si + 2] = z + a.d_xyz[3 * si + 2];
  } else {
    const float4* T = (const float4*)(a.d_xyz + 16 * si);
    float m[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const float4 t = T[r];
      m[r] = t.x * x + t.y * y + t.z * z + t.w;
    }
    means3D[3 * si] = m[0] / m[3];
    means3D[3 * si + 1] = m[1] / m[3];
    means3D[3 * si + 2] = m[2] / m[3];
  }
  opac[si] = 1.0f / (1.0f + expf(-a.opacity[si]));
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float s = expf(a.scaling[3 * si + k]);
    if (a.d_scaling) s += a.d_scaling[3 * si + k];
    scales[3 * si + k] = s;
  }
  const float4 q = *(const float4*)(a.rotation + 4 * si);
  const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
  float4 o = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
  if (a.d_rotation) {
    const float4 d = *(const float4*)(a.d_rotation + 4 * si);
    o.x += d.x; o.y += d.y; o.z += d.z; o.w += d.w;
  }
  *(float4*)(rots + 4 * si) = o;
}

// Cotangents of (means3D, opacity, scales, rotations) -> cotangents of the raw parameters.  The
// cotangents of the additive deltas (d_xyz non-6dof, d_scaling, d_rotation) are the incoming tensors
// themselves; only the 6-DoF transform gets its own [N,4,4] gradient here.
__global__ void __launch_bounds__(256)
prepare_bwd_kernel(PrepArgs a, const float* __restrict__ g_means3D, const float* __restrict__ g_opac,
                   const float* __restrict__ g_scales, const float* __restrict__ g_rots,
                   float* __restrict__ g_xyz, float* __restrict__ g_T, float* __restrict__ g_opacity,
                   float* __restrict__ g_scaling, float* __restrict__ g_rotation, const bool acc) {
  // acc: the four parameter cotangents are added to what g_xyz / g_opacity / g_scaling / g_rotation hold (the
  // parameters' .grad tensors) instead of overwriting them; each thread owns its elements, so plain read-add-write.
  // (__fadd_rn: the sum is never contracted into the product that feeds it, so the result is bit-identical to
  //  writing the cotangent out and letting autograd add it)
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.N) return;
  const size_t si = (size_t)i;
  if (!a.is_6dof && acc && g_xyz && g_means3D) {
#pragma unroll
    for (int k = 0; k < 3; k++) g_xyz[3 * si + k] = __fadd_rn(g_xyz[3 * si + k], g_means3D[3 * si + k]);
  }
  if (a.is_6dof && (g_xyz || g_T)) {
    const float x = a.xyz[3 * si], y = a.xyz[3 * si + 1], z = a.xyz[3 * si + 2];
    const float4* T = (const float4*)(a.d_xyz + 16 * si);
    float4 t[4];
    float m[4];
#pragma unroll
    for (int r = 0; r < 4; r++) { t[r] = T[r]; m[r] = t[r].x * x + t[r].y * y + t[r].z * z + t[r].w; }
    const float iw = 1.0f / m[3];
    float gm[4] = {0.f, 0.f, 0.f, 0.f};
    if (g_means3D) {
      gm[0] = g_means3D[3 * si] * iw; gm[1] = g_means3D[3 * si + 1] * iw; gm[2] = g_means3D[3 * si + 2] * iw;
      gm[3] = -(gm[0] * m[0] + gm[1] * m[1] + gm[2] * m[2]) * iw;
    }
    if (g_T) {
      float4* o = (float4*)(g_T + 16 * si);
#pragma unroll
      for (int r = 0; r < 4; r++) o[r] = make_float4(gm[r] * x, gm[r] * y, gm[r] * z, gm[r]);
    }
    if (g_xyz) {
      const float gx = gm[0] * t[0].x + gm[1] * t[1].x + gm[2] * t[2].x + gm[3] * t[3].x;
      const float gy = gm[0] * t[0].y + gm[1] * t[1].y + gm[2] * t[2].y + gm[3] * t[3].y;
      const float gz = gm[0] * t[0].z + gm[1] * t[1].z + gm[2] * t[2].z + gm[3] * t[3].z;
      g_xyz[3 * si] = acc ? __fadd_rn(g_xyz[3 * si], gx) : gx;
      g_xyz[3 * si + 1] = acc ? __fadd_rn(g_xyz[3 * si + 1], gy) : gy;
      g_xyz[3 * si + 2] = acc ? __fadd_rn(g_xyz[3 * si + 2], gz) : gz;
    }
  }
  if (g_opacity) {
    const float s = 1.0f / (1.0f + expf(-a.opacity[si]));
    const float v = g_opac ? g_opac[si] * (1.0f - s) * s : 0.f;
    g_opacity[si] = acc ? __fadd_rn(g_opacity[si], v) : v;
  }
  if (g_scaling) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float v = g_scales ? g_scales[3 * si + k] * expf(a.scaling[3 * si + k]) : 0.f;
      g_scaling[3 * si + k] = acc ? __fadd_rn(g_scaling[3 * si + k], v) : v;
    }
  }
  if (g_rotation) {
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g_rots) {
      const float4 q = *(const float4*)(a.rotation + 4 * si);
      const float4 g = *(const float4*)(g_rots + 4 * si);
      const float nn = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
      if (nn > 1e-12f) {
        const float in = 1.0f / nn;
        const float dot = (g.x * q.x + g.y * q.y + g.z * q.z + g.w * q.w) * in * in;
        o = make_float4((g.x - q.x * dot) * in, (g.y - q.y * dot) * in, (g.z - q.z * dot) * in, (g.w - q.w * dot) * in);
      } else {  // clamped denominator: y = q / eps
        o = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
      }
    }
    if (acc) {
      const float4 h = *(const float4*)(g_rotation + 4 * si);
      o.x = __fadd_rn(h.x, o.x); o.y = __fadd_rn(h.y, o.y); o.z = __fadd_rn(h.z, o.z); o.w = __fadd_rn(h.w, o.w);
    }
    *(float4*)(g_rotation + 4 * si) = o;
  }
}

}  // namespace d3g

using namespace d3g;

static int fill_args(PrepArgs& a, int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                     const float* scaling, const float* d_scaling, const float* rotation,
                     const float* d_rotation, const float* opacity) {
  if (N < 0) return set_err(D3G_ERR_INVALID, "negative N");
  if (N > 0 && (!xyz || !d_xyz || !scaling || !rotation || !opacity))
    return set_err(D3G_ERR_INVALID, "NULL tensor");
  a.N = N; a.is_6dof = is_6dof; a.xyz = xyz; a.d_xyz = d_xyz; a.scaling = scaling; a.d_scaling = d_scaling;
  a.rotation = rotation; a.d_rotation = d_rotation; a.opacity = opacity;
  return 0;
}

extern "C" int d3g_gaussians_prepare(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                                     const float* scaling, const float* d_scaling, const float* rotation,
                                     const float* d_rotation, const float* opacity, float* means3D,
                                     float* opacity_out, float* scales_out, float* rotations_out,
                                     void* stream_) {
  PrepArgs a;
  if (int e = fill_args(a, N, is_6dof, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity)) return e;
  if (N > 0 && (!means3D || !opacity_out || !scales_out || !rotations_out))
    return set_err(D3G_ERR_INVALID, "NULL output");
  if (N == 0) return 0;
  cudaStream_t stream = (cudaStream_t)stream_;
  prepare_fwd_kernel<<<(N + 255) / 256, 256, 0, stream>>>(a, means3D, opacity_out, scales_out, rotations_out);
  D3G_LAUNCH_CHECK(0, stream);
  return 0;
}

static int prepare_backward_impl(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                                 const float* scaling, const float* rotation, const float* opacity,
                                 const float* g_means3D, const float* g_opacity_out, const float* g_scales_out,
                                 const float* g_rotations_out, float* g_xyz, float* g_dxyz_6dof, float* g_opacity,
                                 float* g_scaling, float* g_rotation, bool acc, void* stream_) {
  PrepArgs a;
  if (int e = fill_args(a, N, is_6dof, xyz, d_xyz, scaling, nullptr, rotation, nullptr, opacity)) return e;
  if (N == 0) return 0;
  cudaStream_t stream = (cudaStream_t)stream_;
  prepare_bwd_kernel<<<(N + 255) / 256, 256, 0, stream>>>(a, g_means3D, g_opacity_out, g_scales_out, g_rotations_out,
                                                          g_xyz, g_dxyz_6dof, g_opacity, g_scaling, g_rotation, acc);
  D3G_LAUNCH_CHECK(0, stream);
  return 0;
}

extern "C" int d3g_gaussians_prepare_backward(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                                              const float* scaling, const float* rotation, const float* opacity,
                                              const float* g_means3D, const float* g_opacity_out,
                                              const float* g_scales_out, const float* g_rotations_out,
                                              float* g_xyz, float* g_dxyz_6dof, float* g_opacity,
                                              float* g_scaling, float* g_rotation, void* stream_) {
  return prepare_backward_impl(N, is_6dof, xyz, d_xyz, scaling, rotation, opacity, g_means3D, g_opacity_out,
                               g_scales_out, g_rotations_out, g_xyz, g_dxyz_6dof, g_opacity, g_scaling, g_rotation,
                               false, stream_);
}

extern "C" int d3g_gaussians_prepare_backward_into(int32_t N, int32_t is_6dof, const float* xyz, const float* d_xyz,
                                                   const float* scaling, const float* rotation, const float* opacity,
                                                   const float* g_means3D, const float* g_opacity_out,
                                                   const float* g_scales_out, const float* g_rotations_out,
                                                   float* acc_xyz, float* g_dxyz_6dof, float* acc_opacity,
                                                   float* acc_scaling, float* acc_rotation, void* stream_) {
  return prepare_backward_impl(N, is_6dof, xyz, d_xyz, scaling, rotation, opacity, g_means3D, g_opacity_out,
                               g_scales_out, g_rotations_out, acc_xyz, g_dxyz_6dof, acc_opacity, acc_scaling,
                               acc_rotation, true, stream_);
}

// ------------------------------------------------------------------------------------------------
// Densification bookkeeping of one rendered frame, fused:
//   max_radii2D[vis] = max(max_radii2D[vis], radii[vis])                         (train.py:238-239)
//   xyz_gradient_accum[vis] += || viewspace_grad[vis, :2] ||;  denom[vis] += 1   (scene/gaussian_model.py:398-401)
// with vis = radii > 0 (gaussian_renderer/__init__.py:131).  One thread per Gaussian, one pass, no boolean-mask
// gathers / scatters (the reference's indexed assignments are ~10 launches and two nonzero() syncs per frame).
// ------------------------------------------------------------------------------------------------
namespace d3g {
__global__ void __launch_bounds__(256)
densification_stats_kernel(int N, const float* __restrict__ vs_grad, const int* __restrict__ radii,
                           float* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radii) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int r = radii[i];
  if (r <= 0) return;
  if (max_radii) max_radii[i] = fmaxf(max_radii[i], (float)r);
  if (vs_grad) {
    const float gx = vs_grad[3 * (size_t)i], gy = vs_grad[3 * (size_t)i + 1];
    accum[i] += sqrtf(gx * gx + gy * gy);
    denom[i] += 1.0f;
  }
}
}  // namespace d3g

extern "C" int d3g_densification_stats(int32_t N, const float* viewspace_grad, const int32_t* radii,
                                       float* xyz_gradient_accum, float* denom, float* max_radii2D, void* stream_) {
  if (N < 0 || (N > 0 && !radii)) return d3g::set_err(D3G_ERR_INVALID, "densification_stats: radii is NULL");
  if (viewspace_grad && (!xyz_gradient_accum || !denom))
    return d3g::set_err(D3G_ERR_INVALID, "densification_stats: accumulators missing");
  if (N == 0) return 0;
  cudaStream_t stream = (cudaStream_t)stream_;
  { d3g::ProfScope _p(d3g::PROF_DENSIFY, stream);
    d3g::densification_stats_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, viewspace_grad, radii, xyz_gradient_accum,
                                                                       denom, max_radii2D);
    D3G_LAUNCH_CHECK(0, stream); }
  return 0;
}
