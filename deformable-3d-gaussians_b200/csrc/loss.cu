// Fused image loss of the training step: L1 + SSIM, forward and gradient.
//
// Replaces utils/loss_utils.py:18-19 (l1_loss) and :42-73 (ssim / _ssim: 11x11 Gaussian window, sigma 1.5,
// zero padding 5, one window per channel, C1 = 0.01^2, C2 = 0.03^2, mean over the map), and the
// combination of train_baseline.py:126-127:  loss = (1 - lambda) * L1 + lambda * (1 - ssim).
// The reference runs 5 grouped conv2d + ~15 elementwise launches forward and as many again in autograd; here
//   ssim_fwd_kernel : one CTA per 32x16 pixel block of one channel stages a 42x26 halo of both images in shared
//                     memory, runs the separable window (5 moments), evaluates the SSIM map, and stores the three
//                     per-pixel partials dS/dmu1, dS/dE[x^2], dS/dE[xy] plus per-block partial sums of S and |x-y|;
//   loss_reduce_kernel : deterministic tree sum of the partials -> {L1 mean, SSIM mean, combined loss};
//   ssim_bwd_kernel : the transposed (= same, symmetric) window applied to the three partial maps, combined with
//                     the L1 sign term into dL/dimage.
// The gradient is produced together with the value (it does not depend on the upstream cotangent except as a
// scalar factor, which the host side applies).
#include "common.cuh"

namespace d3g {

constexpr int kTW = 32, kTH = 16;  // pixel block of one CTA (256 threads, two pixels each)
constexpr int kHalo = 5;           // window 11
constexpr int kSW = kTW + 2 * kHalo, kSH = kTH + 2 * kHalo;  // 42 x 26 staged halo
constexpr int kLThreads = 256;

__constant__ float c_win[11];
static bool g_win_ready[64] = {};   // per device: constant memory is a per-device object

// utils/loss_utils.py:31-33: exp(-(x - 5)^2 / (2 * 1.5^2)) as Python floats, stored as float32, divided by
// their float32 sum.
static void make_window(float* w) {
  float s = 0.f;
  for (int x = 0; x < 11; x++) {
    w[x] = (float)exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5));
    s += w[x];
  }
  for (int x = 0; x < 11; x++) w[x] = w[x] / s;
}

struct LossArgs {
  int C, H, W;
  const float* img;
  const float* gt;
  float* dmu;    // dS/dmu1
  float* d11;    // dS/dE[x^2]
  float* d12;    // dS/dE[xy]
  float* part;   // [blocks][2]: sum S, sum |x-y|
};

__global__ void __launch_bounds__(kLThreads)
ssim_fwd_kernel(LossArgs a) {
  __shared__ float s1[kSH][kSW + 1];
  __shared__ float s2[kSH][kSW + 1];
  __shared__ float h[5][kSH][kTW + 1];   // horizontally filtered moments
  __shared__ float red[2][kLThreads / 32];
  const int c = blockIdx.z;
  const int x0 = blockIdx.x * kTW, y0 = blockIdx.y * kTH;
  const size_t plane = (size_t)c * a.H * a.W;
  for (int i = threadIdx.x; i < kSH * kSW; i += kLThreads) {
    const int ly = i / kSW, lx = i % kSW;
    const int gy = y0 + ly - kHalo, gx = x0 + lx - kHalo;
    float v1 = 0.f, v2 = 0.f;   // zero padding, as F.conv2d(padding=5)
    if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
      v1 = a.img[plane + (size_t)gy * a.W + gx];
      v2 = a.gt[plane + (size_t)gy * a.W + gx];
    }
    s1[ly][lx] = v1;
    s2[ly][lx] = v2;
  }
  __syncthreads();
  // horizontal pass: 26 rows x 16 columns, 5 moments each
  for (int i = threadIdx.x; i < kSH * kTW; i += kLThreads) {
    const int ly = i / kTW, lx = i % kTW;
    float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
      const float w = c_win[k], p = s1[ly][lx + k], q = s2[ly][lx + k];
      m1 += w * p; m2 += w * q; e11 += w * p * p; e22 += w * q * q; e12 += w * p * q;
    }
    h[0][ly][lx] = m1; h[1][ly][lx] = m2; h[2][ly][lx] = e11; h[3][ly][lx] = e22; h[4][ly][lx] = e12;
  }
  __syncthreads();
  const int tx = threadIdx.x % kTW;
  float S = 0.f, l1 = 0.f;
  // a thread owns two vertically adjacent pixels: their windows share 10 of 11 rows, so the column is read once
  const int ty0 = (threadIdx.x / kTW) * 2;
  float vm[2][5];
#pragma unroll
  for (int r = 0; r < 2; r++)
#pragma unroll
    for (int m = 0; m < 5; m++) vm[r][m] = 0.f;
#pragma unroll
  for (int k = 0; k < 12; k++) {
#pragma unroll
    for (int m = 0; m < 5; m++) {
      const float v = h[m][ty0 + k][tx];
      if (k < 11) vm[0][m] += c_win[k] * v;
      if (k >= 1) vm[1][m] += c_win[k - 1] * v;
    }
  }
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int ty = ty0 + r;
    const float mu1 = vm[r][0], mu2 = vm[r][1], e11 = vm[r][2], e22 = vm[r][3], e12 = vm[r][4];
    const int gx = x0 + tx, gy = y0 + ty;
    if (gx < a.W && gy < a.H) {
      const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
      const float sig1 = e11 - mu1_sq, sig2 = e22 - mu2_sq, sig12 = e12 - mu12;
      const float A1 = 2.f * mu12 + C1, A2 = 2.f * sig12 + C2, B1 = mu1_sq + mu2_sq + C1, B2 = sig1 + sig2 + C2;
      const float inv = 1.0f / (B1 * B2);
      const float Sp = A1 * A2 * inv;
      S += Sp;
      const size_t o = plane + (size_t)gy * a.W + gx;
      // partials of S w.r.t. the three moments that depend on img1 (sigma1^2 = E11 - mu1^2, sigma12 = E12 - mu1 mu2)
      a.dmu[o] = 2.f * mu2 * (A2 - A1) * inv - Sp * 2.f * mu1 * (1.0f / B1 - 1.0f / B2);
      a.d11[o] = -Sp / B2;
      a.d12[o] = 2.f * A1 * inv;
      l1 += fabsf(s1[ty + kHalo][tx + kHalo] - s2[ty + kHalo][tx + kHalo]);
    }
  }
  // block partial sums (fixed order: deterministic)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    S += __shfl_xor_sync(0xffffffffu, S, o);
    l1 += __shfl_xor_sync(0xffffffffu, l1, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = S; red[1][threadIdx.x >> 5] = l1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ts = 0.f, tl = 0.f;
    for (int w = 0; w < kLThreads / 32; w++) { ts += red[0][w]; tl += red[1][w]; }
    const size_t b = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    a.part[2 * b] = ts;
    a.part[2 * b + 1] = tl;
  }
}

// out[0] = mean |x - y|, out[1] = mean SSIM, out[2] = w_l1 * out[0] + w_ssim * out[1] + bias
__global__ void __launch_bounds__(1024)
loss_reduce_kernel(const float* __restrict__ part, int blocks, float inv_n, float w_l1, float w_ssim, float bias,
                   float* __restrict__ out) {
  __shared__ double rs[32], rl[32];
  double s = 0.0, l = 0.0;
  for (int i = threadIdx.x; i < blocks; i += 1024) { s += part[2 * i]; l += part[2 * i + 1]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    l += __shfl_xor_sync(0xffffffffu, l, o);
  }
  if ((threadIdx.x & 31) == 0) { rs[threadIdx.x >> 5] = s; rl[threadIdx.x >> 5] = l; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, tl = 0.0;
    for (int w = 0; w < 32; w++) { ts += rs[w]; tl += rl[w]; }
    const float l1 = (float)(tl * inv_n), ss = (float)(ts * inv_n);
    out[0] = l1;
    out[1] = ss;
    out[2] = w_l1 * l1 + w_ssim * ss + bias;
  }
}

// dL/dimg = w_l1/n * sign(x - y) + w_ssim/n * [ win*dmu + 2 x (win*d11) + y (win*d12) ]
__global__ void __launch_bounds__(kLThreads)
ssim_bwd_kernel(LossArgs a, float w_l1_n, float w_ssim_n, const float* __restrict__ upstream, float* __restrict__ dimg) {
  __shared__ float s[3][kSH][kSW + 1];
  __shared__ float h[3][kSH][kTW + 1];
  const int c = blockIdx.z;
  const int x0 = blockIdx.x * kTW, y0 = blockIdx.y * kTH;
  const size_t plane = (size_t)c * a.H * a.W;
  if (upstream) {   // cotangent of the scalar loss (device scalar), folded in here instead of a separate pass
    const float u = __ldg(upstream);
    w_l1_n *= u;
    w_ssim_n *= u;
  }
  if (w_ssim_n != 0.f) {
    for (int i = threadIdx.x; i < kSH * kSW; i += kLThreads) {
      const int ly = i / kSW, lx = i % kSW;
      const int gy = y0 + ly - kHalo, gx = x0 + lx - kHalo;
      float v0 = 0.f, v1 = 0.f, v2 = 0.f;
      if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W) {
        const size_t o = plane + (size_t)gy * a.W + gx;
        v0 = a.dmu[o]; v1 = a.d11[o]; v2 = a.d12[o];
      }
      s[0][ly][lx] = v0; s[1][ly][lx] = v1; s[2][ly][lx] = v2;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kSH * kTW; i += kLThreads) {
      const int ly = i / kTW, lx = i % kTW;
      float r0 = 0.f, r1 = 0.f, r2 = 0.f;
#pragma unroll
      for (int k = 0; k < 11; k++) {
        const float w = c_win[k];
        r0 += w * s[0][ly][lx + k]; r1 += w * s[1][ly][lx + k]; r2 += w * s[2][ly][lx + k];
      }
      h[0][ly][lx] = r0; h[1][ly][lx] = r1; h[2][ly][lx] = r2;
    }
    __syncthreads();
  }
  const int tx = threadIdx.x % kTW;
  const int ty0 = (threadIdx.x / kTW) * 2;
  float vr[2][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
  if (w_ssim_n != 0.f) {
#pragma unroll
    for (int k = 0; k < 12; k++) {
#pragma unroll
      for (int m = 0; m < 3; m++) {
        const float v = h[m][ty0 + k][tx];
        if (k < 11) vr[0][m] += c_win[k] * v;
        if (k >= 1) vr[1][m] += c_win[k - 1] * v;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int ty = ty0 + r;
    const int gx = x0 + tx, gy = y0 + ty;
    if (gx >= a.W || gy >= a.H) continue;
    const size_t o = plane + (size_t)gy * a.W + gx;
    const float x = a.img[o], y = a.gt[o];
    float g = 0.f;
    if (w_ssim_n != 0.f) g = w_ssim_n * (vr[r][0] + 2.f * x * vr[r][1] + y * vr[r][2]);
    const float d = x - y;
    g += w_l1_n * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));   // torch.sign: 0 at 0
    dimg[o] = g;
  }
}

}  // namespace d3g

using namespace d3g;

extern "C" {

int d3g_image_loss_workspace_bytes(int32_t C, int32_t H, int32_t W, size_t* bytes) {
  if (C < 1 || H < 1 || W < 1 || !bytes) return set_err(D3G_ERR_INVALID, "bad image_loss shape");
  const size_t blocks = (size_t)C * ((H + kTH - 1) / kTH) * ((W + kTW - 1) / kTW);
  *bytes = align_up(3 * (size_t)C * H * W * sizeof(float), 256) + align_up(2 * blocks * sizeof(float), 256);
  return 0;
}

int d3g_image_loss(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float w_l1, float w_ssim,
                   float bias, void* workspace, float* out3, float* dL_dimg, void* stream_) {
  if (C < 1 || H < 1 || W < 1) return set_err(D3G_ERR_INVALID, "bad image_loss shape");
  if (!img || !gt || !workspace || !out3) return set_err(D3G_ERR_INVALID, "NULL argument");
  cudaStream_t stream = (cudaStream_t)stream_;
  int dev = 0;
  D3G_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !g_win_ready[dev]) {
    float w[11];
    make_window(w);
    D3G_CUDA_TRY(cudaMemcpyToSymbolAsync(c_win, w, sizeof(w), 0, cudaMemcpyHostToDevice, stream));
    D3G_CUDA_TRY(cudaStreamSynchronize(stream));   // w is a stack buffer; once per device
    if (dev >= 0 && dev < 64) g_win_ready[dev] = true;
  }
  const dim3 grid((W + kTW - 1) / kTW, (H + kTH - 1) / kTH, C);
  const size_t blocks = (size_t)grid.x * grid.y * grid.z;
  if (blocks > 0x7fffffffull) return set_err(D3G_ERR_CAPACITY, "image too large");
  const size_t n = (size_t)C * H * W;
  LossArgs a;
  a.C = C; a.H = H; a.W = W; a.img = img; a.gt = gt;
  a.dmu = (float*)workspace; a.d11 = a.dmu + n; a.d12 = a.d11 + n;
  a.part = (float*)((uint8_t*)workspace + align_up(3 * n * sizeof(float), 256));
  { ProfScope _p(PROF_LOSS_FWD, stream);
    ssim_fwd_kernel<<<grid, kLThreads, 0, stream>>>(a);
    D3G_LAUNCH_CHECK(0, stream);
    loss_reduce_kernel<<<1, 1024, 0, stream>>>(a.part, (int)blocks, (float)(1.0 / (double)n), w_l1, w_ssim, bias, out3);
    D3G_LAUNCH_CHECK(0, stream); }
  if (dL_dimg) return d3g_image_loss_backward(C, H, W, img, gt, w_l1, w_ssim, workspace, nullptr, dL_dimg, stream_);
  return 0;
}

int d3g_image_loss_backward(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float w_l1,
                            float w_ssim, const void* workspace, const float* upstream, float* dL_dimg,
                            void* stream_) {
  if (C < 1 || H < 1 || W < 1) return set_err(D3G_ERR_INVALID, "bad image_loss shape");
  if (!img || !gt || !workspace || !dL_dimg) return set_err(D3G_ERR_INVALID, "NULL argument");
  cudaStream_t stream = (cudaStream_t)stream_;
  const dim3 grid((W + kTW - 1) / kTW, (H + kTH - 1) / kTH, C);
  const size_t n = (size_t)C * H * W;
  LossArgs a;
  a.C = C; a.H = H; a.W = W; a.img = img; a.gt = gt;
  a.dmu = (float*)workspace; a.d11 = a.dmu + n; a.d12 = a.d11 + n;
  a.part = nullptr;
  { ProfScope _p(PROF_LOSS_BWD, stream);
    ssim_bwd_kernel<<<grid, kLThreads, 0, stream>>>(a, (float)(w_l1 / (double)n), (float)(w_ssim / (double)n), upstream,
                                                    dL_dimg);
    D3G_LAUNCH_CHECK(0, stream); }
  return 0;
}

// ------------------------------------------------------------------------------------------------
// d3g_image_loss_static: the same loss + gradient for callers that may not allocate (no counterpart in the reference).
// The workspace d3g_image_loss takes from its caller (three per-pixel
// partial maps + the per-block partial sums) lives in a file-scope device array instead, so the call performs no
// cudaMalloc / cudaFree and asks nothing of the caller's allocator.  Every GPU that loads libd3g.so holds the array
// (24.1 MiB) until the process exits.
//
// In host code `g_loss_scratch` is only a host-side stand-in for the symbol, never a device address: the device
// address is looked up once per device, with that device current, through cudaGetSymbolAddress.
// ------------------------------------------------------------------------------------------------
}  // extern "C"

namespace d3g {
constexpr size_t kStaticLossMaxElems = (size_t)1 << 21;    // C*H*W of one call (3 x 800 x 800 = 1,920,000 fits)
constexpr size_t kStaticLossMaxBlocks = (size_t)1 << 14;   // C * ceil(H/16) * ceil(W/32) of one call
constexpr size_t kStaticLossBytes = 3 * kStaticLossMaxElems * sizeof(float)        // dS/dmu1, dS/dE[x^2], dS/dE[xy]
                                    + 2 * kStaticLossMaxBlocks * sizeof(float);   // per-block {sum S, sum |x-y|}
static __device__ __align__(256) unsigned char g_loss_scratch[kStaticLossBytes];
static void* g_loss_scratch_ptr[64] = {};   // per device: the symbol's address on that device
}  // namespace d3g

extern "C" {

int d3g_image_loss_static_limits(size_t* max_elements, size_t* max_blocks, size_t* scratch_bytes) {
  if (max_elements) *max_elements = kStaticLossMaxElems;
  if (max_blocks) *max_blocks = kStaticLossMaxBlocks;
  if (scratch_bytes) *scratch_bytes = kStaticLossBytes;
  return 0;
}

int d3g_image_loss_static(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float w_l1,
                          float w_ssim, float bias, float* out3, float* dL_dimg, void* stream_) {
  if (C < 1 || H < 1 || W < 1) return set_err(D3G_ERR_INVALID, "bad image_loss shape");
  if (!img || !gt || !out3) return set_err(D3G_ERR_INVALID, "NULL argument");
  const size_t n = (size_t)C * H * W;
  const size_t blocks = (size_t)C * ((H + kTH - 1) / kTH) * ((W + kTW - 1) / kTW);
  if (n > kStaticLossMaxElems)
    return set_err(D3G_ERR_CAPACITY, "d3g_image_loss_static: C*H*W = %zu exceeds the static scratch limit of %zu elements "
                   "(use d3g_image_loss with a caller-provided workspace)", n, (size_t)kStaticLossMaxElems);
  if (blocks > kStaticLossMaxBlocks)
    return set_err(D3G_ERR_CAPACITY, "d3g_image_loss_static: %zu 32x16 pixel blocks exceed the static scratch limit of %zu "
                   "(use d3g_image_loss with a caller-provided workspace)", blocks, (size_t)kStaticLossMaxBlocks);
  int dev = 0;
  D3G_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return set_err(D3G_ERR_UNSUPPORTED, "d3g_image_loss_static: device index %d above 63", dev);
  if (!g_loss_scratch_ptr[dev]) {   // set-up, once per device, with that device current
    void* ptr = nullptr;
    D3G_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_loss_scratch));
    if (!ptr) return set_err(D3G_ERR_CUDA, "d3g_image_loss_static: cudaGetSymbolAddress returned NULL");
    g_loss_scratch_ptr[dev] = ptr;
  }
  // Same kernels, same workspace layout as d3g_image_loss (which sizes the layout from this call's n and blocks, both
  // within the array by the checks above).  Nothing in the array has to start at zero: every word a kernel reads
  // (the partial maps at in-image pixels, the block sums) is written earlier in the same call, so the contents the
  // previous call left behind are never seen.
  return d3g_image_loss(C, H, W, img, gt, w_l1, w_ssim, bias, g_loss_scratch_ptr[dev], out3, dL_dimg, stream_);
}

}  // extern "C"
