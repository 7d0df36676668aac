// Fused multi-tensor Adam step: the optimizer step that follows the hot path every
// iteration — `GaussianModel.optimizer.step()` over six parameter groups (scene/gaussian_model.py:127-136) and
// `DeformModel.optimizer.step()` over the MLP parameters (scene/deform_model.py:275-295,326-337), both
// torch.optim.Adam(lr per group, betas (0.9, 0.999), eps 1e-15, optional weight_decay, no amsgrad).
// One launch updates every tensor: the segment table travels as a kernel argument (no H2D copy), elements are
// processed 4 at a time where the four arrays are 16-byte aligned.  Arithmetic follows torch's single-tensor
// reference (_single_tensor_adam): g += wd*p; m = lerp(m, g, 1-b1); v = v*b2 + (1-b2)*g*g;
// p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps).
#include "common.cuh"

namespace d3g {

constexpr int kAdamMaxSeg = 48;
constexpr int kAdamThreads = 256;
constexpr int kAdamPerBlock = kAdamThreads * 4 * 4;   // elements per CTA (four float4 per thread)

struct AdamSeg {
  float* p; const float* g; float* m; float* v;
  int64_t n;
  float step_size;      // lr / (1 - beta1^t)
  float bc2_sqrt;       // sqrt(1 - beta2^t)
  float weight_decay;
  int first_block;      // prefix sum of CTAs
};
struct AdamArgs {
  AdamSeg seg[kAdamMaxSeg];
  int n_seg;
  float beta1, beta2, eps;
  float om_beta1, om_beta2;   // 1 - beta computed in double on the host, as torch's Python scalars are
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamSeg& s, const AdamArgs& a) {
  const float b2 = a.beta2, eps = a.eps;
  if (s.weight_decay != 0.f) g = fmaf(s.weight_decay, p, g);
  m = m + a.om_beta1 * (g - m);
  v = v * b2 + a.om_beta2 * g * g;
  const float denom = sqrtf(v) / s.bc2_sqrt + eps;   // torch divides (exp_avg_sq.sqrt() / bias_correction2_sqrt)
  p = p - s.step_size * (m / denom);
}

__global__ void __launch_bounds__(kAdamThreads)
adam_kernel(const __grid_constant__ AdamArgs a) {
  // which segment does this CTA belong to (<= 48 segments: linear scan over the prefix sums)
  int si = 0;
  while (si + 1 < a.n_seg && (int)blockIdx.x >= a.seg[si + 1].first_block) si++;
  const AdamSeg& s = a.seg[si];
  const int64_t base = (int64_t)((int)blockIdx.x - s.first_block) * kAdamPerBlock;
  const bool vec = ((((uintptr_t)s.p | (uintptr_t)s.g | (uintptr_t)s.m | (uintptr_t)s.v) & 15) == 0);
#pragma unroll
  for (int r = 0; r < 4; r++) {
    const int64_t i = base + ((int64_t)r * kAdamThreads + threadIdx.x) * 4;
    if (i >= s.n) break;
    if (vec && i + 4 <= s.n) {
      float4 p = *(float4*)(s.p + i), m = *(float4*)(s.m + i), v = *(float4*)(s.v + i);
      const float4 g = *(const float4*)(s.g + i);
      adam_one(p.x, g.x, m.x, v.x, s, a);
      adam_one(p.y, g.y, m.y, v.y, s, a);
      adam_one(p.z, g.z, m.z, v.z, s, a);
      adam_one(p.w, g.w, m.w, v.w, s, a);
      *(float4*)(s.p + i) = p; *(float4*)(s.m + i) = m; *(float4*)(s.v + i) = v;
    } else {
      for (int k = 0; k < 4 && i + k < s.n; k++) {
        float p = s.p[i + k], m = s.m[i + k], v = s.v[i + k];
        adam_one(p, s.g[i + k], m, v, s, a);
        s.p[i + k] = p; s.m[i + k] = m; s.v[i + k] = v;
      }
    }
  }
}

}  // namespace d3g

using namespace d3g;

extern "C" int d3g_adam_step(int32_t n_tensors, const d3g_adam_tensor* t, double beta1_d, double beta2_d, double eps_d,
                             void* stream_) {
  if (n_tensors < 0 || (n_tensors > 0 && !t)) return set_err(D3G_ERR_INVALID, "bad adam arguments");
  cudaStream_t stream = (cudaStream_t)stream_;
  // betas / eps cross the ABI as the Python doubles torch itself uses: 1 - beta and the bias corrections are taken
  // from them in double, like torch.optim.Adam's scalar path
  if (!(beta1_d >= 0.0 && beta1_d < 1.0) || !(beta2_d >= 0.0 && beta2_d < 1.0))
    return set_err(D3G_ERR_INVALID, "adam: betas must lie in [0, 1)");
  const float beta1 = (float)beta1_d, beta2 = (float)beta2_d, eps = (float)eps_d;
  int done = 0;
  while (done < n_tensors) {   // more than 48 tensors: several launches
    AdamArgs a;
    a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    a.om_beta1 = (float)(1.0 - (double)beta1_d); a.om_beta2 = (float)(1.0 - (double)beta2_d);
    int blocks = 0, k = 0;
    for (; k < kAdamMaxSeg && done + k < n_tensors; k++) {
      const d3g_adam_tensor& x = t[done + k];
      if (x.numel < 0 || (x.numel > 0 && (!x.param || !x.grad || !x.exp_avg || !x.exp_avg_sq)) || x.step < 1)
        return set_err(D3G_ERR_INVALID, "adam tensor %d: NULL pointer, negative size or step < 1", done + k);
      AdamSeg& s = a.seg[k];
      s.p = x.param; s.g = x.grad; s.m = x.exp_avg; s.v = x.exp_avg_sq; s.n = x.numel;
      const double bc1 = 1.0 - pow(beta1_d, (double)x.step), bc2 = 1.0 - pow(beta2_d, (double)x.step);
      s.step_size = (float)((double)x.lr / bc1);
      s.bc2_sqrt = (float)sqrt(bc2);
      s.weight_decay = x.weight_decay;
      s.first_block = blocks;
      const int64_t nb = (x.numel + kAdamPerBlock - 1) / kAdamPerBlock;
      if (blocks + nb > 0x7fffffff) return set_err(D3G_ERR_CAPACITY, "adam: too many elements for one launch");
      blocks += (int)nb;
    }
    a.n_seg = k;
    if (blocks > 0) {
      ProfScope _p(PROF_ADAM, stream);
      adam_kernel<<<blocks, kAdamThreads, 0, stream>>>(a);
      D3G_LAUNCH_CHECK(0, stream);
    }
    done += k;
  }
  return 0;
}
