// Tile-based differentiable Gaussian rasterizer for sm_100a — forward and backward.
//
// Replaces the CUDA extension the reference calls at gaussian_renderer/__init__.py:115-124
// (diff_gaussian_rasterization; source absent from the reference tree).  Algorithm = SURVEY.md
// Appendix A; the floating-point operation order of every stage that feeds an integer artefact
// (radii, tiles_touched, keys, ranges) and of the forward blend is pinned to the oracle
// (oracle/raster_ref_impl.h) with explicitly rounded intrinsics, so those results are bit-exact.
//
// HBM layout (all SoA, 16-byte records so every gather is one LDG.128):
//   geom:  header{R,overflow} | means2D float2[N] | conic_opacity float4[N] | rgbd float4[N]
//          | cov3D float[6N] | tiles_touched u32[N] | offsets u32[N] | clamped u8[N]
//          | backward accumulators 3 x float4[N] | scan scratch
//   bin:   keys u64[cap] x2 | values u32[cap] x2 | radix-sort scratch
//   img:   ranges uint2[tiles] | final_T f32[P] | n_contrib u32[P]
#include <cub/cub.cuh>

#include "common.cuh"

namespace d3g {

static constexpr int kTile = 16;
static constexpr int kBlock = kTile * kTile;
static constexpr int kTileCtrStride = 32;   // uint32 words between the counters of consecutive tiles (128 B)

struct GeomView {
  uint32_t* header;  // [0]=R [1]=overflow
  float2* means2D;
  float4* conic_opacity;
  float4* rgbd;
  float* cov3D;
  uint32_t* tiles_touched;
  uint32_t* offsets;
  uint8_t* clamped;
  float4* acc_mean;   // (dmean2D.x, dmean2D.y, densify.x, densify.y)
  float4* acc_conic;  // (dconic.xx, dconic.xy, dconic.yy, dopacity)
  float4* acc_color;  // (dr, dg, db, dz)
  void* scan_temp;
  size_t scan_temp_bytes;
  size_t total;
};

static size_t scan_temp_bytes_for(int N) {
  size_t b = 0;
  cub::DeviceScan::InclusiveSum(nullptr, b, (uint32_t*)nullptr, (uint32_t*)nullptr, N > 0 ? N : 1);
  return b;
}

static GeomView geom_view(void* base, int N) {
  GeomView v;
  size_t n = (size_t)(N > 0 ? N : 1);
  char* p = (char*)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { char* r = p + off; off += align_up(bytes, 256); return r; };
  v.header = (uint32_t*)take(256);
  v.means2D = (float2*)take(n * 8);
  v.conic_opacity = (float4*)take(n * 16);
  v.rgbd = (float4*)take(n * 16);
  v.cov3D = (float*)take(n * 24);
  v.tiles_touched = (uint32_t*)take(n * 4);
  v.offsets = (uint32_t*)take(n * 4);
  v.clamped = (uint8_t*)take(n);
  v.acc_mean = (float4*)take(n * 16);
  v.acc_conic = (float4*)take(n * 16);
  v.acc_color = (float4*)take(n * 16);
  v.scan_temp_bytes = scan_temp_bytes_for(N);
  v.scan_temp = take(v.scan_temp_bytes + 256);
  v.total = off;
  return v;
}

struct BinView {
  uint64_t* keys_unsorted;
  uint64_t* keys_sorted;
  uint32_t* vals_unsorted;
  uint32_t* point_list;
  void* sort_temp;
  size_t sort_temp_bytes;
  size_t total;
};

static size_t sort_temp_bytes_for(uint64_t cap) {
  size_t b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, b, (uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (uint32_t*)nullptr, (uint32_t*)nullptr, (int)(cap > 0 ? cap : 1), 0, 64);
  return b;
}

static BinView bin_view(void* base, uint64_t cap) {
  BinView v;
  size_t n = (size_t)(cap > 0 ? cap : 1);
  char* p = (char*)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { char* r = p + off; off += align_up(bytes, 256); return r; };
  v.keys_unsorted = (uint64_t*)take(n * 8);
  v.keys_sorted = (uint64_t*)take(n * 8);
  v.vals_unsorted = (uint32_t*)take(n * 4);
  v.point_list = (uint32_t*)take(n * 4);
  v.sort_temp_bytes = sort_temp_bytes_for(cap);
  v.sort_temp = take(v.sort_temp_bytes + 256);
  v.total = off;
  return v;
}

struct ImgView {
  uint2* ranges;
  float* final_T;
  uint32_t* n_contrib;
  uint32_t* tile_count;    // tile binning: pairs per tile
  uint32_t* tile_cursor;   // tile binning: write cursor of each tile's segment
  size_t total;
};

static ImgView img_view(void* base, int H, int W) {
  ImgView v;
  size_t tiles = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
  size_t P = (size_t)H * W;
  char* p = (char*)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { char* r = p + off; off += align_up(bytes, 256); return r; };
  v.ranges = (uint2*)take((tiles ? tiles : 1) * 8);
  v.final_T = (float*)take((P ? P : 1) * 4);
  v.n_contrib = (uint32_t*)take((P ? P : 1) * 4);
  // one 128-byte line per tile ([0] = count, [1] = cursor): the per-pair atomics of neighbouring tiles then land on
  // different L2 lines instead of serialising on 80 hot ones (measured 90 us -> see DESIGN.md for count / scatter)
  v.tile_count = (uint32_t*)take((tiles ? tiles : 1) * kTileCtrStride * 4);
  v.tile_cursor = v.tile_count + 1;
  v.total = off;
  return v;
}

// ================================================================================================
// device helpers pinned to the oracle's op order
// ================================================================================================
__device__ __forceinline__ void xform4x3(const float* __restrict__ m, const float p[3], float o[3]) {
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float t = fm(__ldg(m + a), p[0]);
    t = ff(__ldg(m + 4 + a), p[1], t);
    t = ff(__ldg(m + 8 + a), p[2], t);
    o[a] = fa(t, __ldg(m + 12 + a));
  }
}
__device__ __forceinline__ void xform4x4(const float* __restrict__ m, const float p[3], float o[4]) {
#pragma unroll
  for (int a = 0; a < 4; a++) {
    float t = fm(__ldg(m + a), p[0]);
    t = ff(__ldg(m + 4 + a), p[1], t);
    t = ff(__ldg(m + 8 + a), p[2], t);
    o[a] = fa(t, __ldg(m + 12 + a));
  }
}

__device__ __forceinline__ float ndc2pix(float v, int S) {
  double t = __dadd_rn((double)v, 1.0);
  double u = __fma_rn(t, (double)S, -1.0);
  return __double2float_rn(__dmul_rn(u, 0.5));
}

__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy, int& mnx,
                                         int& mny, int& mxx, int& mxy) {
  float r = (float)radius;
  int a = __float2int_rz(fdv(fs(px, r), 16.0f));
  int b = __float2int_rz(fdv(fs(py, r), 16.0f));
  int c = __float2int_rz(fdv(fs(fa(fa(px, r), 16.0f), 1.0f), 16.0f));
  int d = __float2int_rz(fdv(fs(fa(fa(py, r), 16.0f), 1.0f), 16.0f));
  mnx = min(gx, max(0, a));
  mny = min(gy, max(0, b));
  mxx = min(gx, max(0, c));
  mxy = min(gy, max(0, d));
}

__device__ __forceinline__ void quat_to_R(const float q[4], float R[9]) {
  float r = q[0], x = q[1], y = q[2], z = q[3];
  float xx = fm(x, x), yy = fm(y, y), zz = fm(z, z);
  float xy = fm(x, y), xz = fm(x, z), yz = fm(y, z), rx = fm(r, x), ry = fm(r, y), rz = fm(r, z);
  R[0] = fs(1.0f, fm(2.0f, fa(yy, zz)));
  R[1] = fm(2.0f, fs(xy, rz));
  R[2] = fm(2.0f, fa(xz, ry));
  R[3] = fm(2.0f, fa(xy, rz));
  R[4] = fs(1.0f, fm(2.0f, fa(xx, zz)));
  R[5] = fm(2.0f, fs(yz, rx));
  R[6] = fm(2.0f, fs(xz, ry));
  R[7] = fm(2.0f, fa(yz, rx));
  R[8] = fs(1.0f, fm(2.0f, fa(xx, yy)));
}

__device__ __forceinline__ void compute_cov3d(const float s_in[3], float mod, const float q[4],
                                              float cov6[6]) {
  float R[9], L[9];
  quat_to_R(q, R);
  float s[3] = {fm(mod, s_in[0]), fm(mod, s_in[1]), fm(mod, s_in[2])};
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int k = 0; k < 3; k++) L[3 * a + k] = fm(R[3 * a + k], s[k]);
  int o = 0;
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int b = a; b < 3; b++) {
      float t = fm(L[3 * a], L[3 * b]);
      t = ff(L[3 * a + 1], L[3 * b + 1], t);
      t = ff(L[3 * a + 2], L[3 * b + 2], t);
      cov6[o++] = t;
    }
}

struct Cov2D {
  float A, B, C;
  float T[6];
  bool clampx, clampy;
  float t[3];
};

__device__ __forceinline__ void compute_cov2d(int W, int H, float tanfovx, float tanfovy,
                                              const float* __restrict__ view, const float pv[3],
                                              const float cov6[6], Cov2D& o) {
  float limx = fm(1.3f, tanfovx), limy = fm(1.3f, tanfovy);
  float tz = pv[2];
  float txtz = fdv(pv[0], tz), tytz = fdv(pv[1], tz);
  o.clampx = (txtz < -limx) || (txtz > limx);
  o.clampy = (tytz < -limy) || (tytz > limy);
  float tx = fm(cmin(limx, cmax(-limx, txtz)), tz);
  float ty = fm(cmin(limy, cmax(-limy, tytz)), tz);
  o.t[0] = tx; o.t[1] = ty; o.t[2] = tz;
  float fx = fdv((float)W, fm(2.0f, tanfovx));
  float fy = fdv((float)H, fm(2.0f, tanfovy));
  float tz2 = fm(tz, tz);
  float J00 = fdv(fx, tz), J02 = fdv(-fm(fx, tx), tz2);
  float J11 = fdv(fy, tz), J12 = fdv(-fm(fy, ty), tz2);
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float w0 = __ldg(view + 4 * k + 0), w1 = __ldg(view + 4 * k + 1), w2 = __ldg(view + 4 * k + 2);
    o.T[k] = ff(J02, w2, fm(J00, w0));
    o.T[3 + k] = ff(J12, w2, fm(J11, w1));
  }
  float S[9] = {cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]};
  float v0[3], v1[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float t = fm(S[3 * a], o.T[0]);
    t = ff(S[3 * a + 1], o.T[1], t);
    v0[a] = ff(S[3 * a + 2], o.T[2], t);
    float u = fm(S[3 * a], o.T[3]);
    u = ff(S[3 * a + 1], o.T[4], u);
    v1[a] = ff(S[3 * a + 2], o.T[5], u);
  }
  float c00 = ff(o.T[2], v0[2], ff(o.T[1], v0[1], fm(o.T[0], v0[0])));
  float c01 = ff(o.T[5], v0[2], ff(o.T[4], v0[1], fm(o.T[3], v0[0])));
  float c11 = ff(o.T[5], v1[2], ff(o.T[4], v1[1], fm(o.T[3], v1[0])));
  o.A = fa(c00, 0.3f);
  o.B = c01;
  o.C = fa(c11, 0.3f);
}

#define D3G_SH_C0 0.28209479177387814f
#define D3G_SH_C1 0.4886025119029199f
__device__ __constant__ float kSHC2[5] = {(float)1.0925484305920792, (float)-1.0925484305920792,
                                          (float)0.31539156525252005, (float)-1.0925484305920792,
                                          (float)0.5462742152960396};
__device__ __constant__ float kSHC3[7] = {(float)-0.5900435899266435, (float)2.890611442640554,
                                          (float)-0.4570457994644658, (float)0.3731763325901154,
                                          (float)-0.4570457994644658, (float)1.445305721320277,
                                          (float)-0.5900435899266435};

__device__ __forceinline__ void sh_basis(int deg, const float d[3], float b[16]) {
  float x = d[0], y = d[1], z = d[2];
  b[0] = D3G_SH_C0;
  if (deg > 0) {
    b[1] = -fm(D3G_SH_C1, y);
    b[2] = fm(D3G_SH_C1, z);
    b[3] = -fm(D3G_SH_C1, x);
    if (deg > 1) {
      float xx = fm(x, x), yy = fm(y, y), zz = fm(z, z), xy = fm(x, y), yz = fm(y, z), xz = fm(x, z);
      b[4] = fm(kSHC2[0], xy);
      b[5] = fm(kSHC2[1], yz);
      b[6] = fm(kSHC2[2], fs(fs(fm(2.0f, zz), xx), yy));
      b[7] = fm(kSHC2[3], xz);
      b[8] = fm(kSHC2[4], fs(xx, yy));
      if (deg > 2) {
        b[9] = fm(fm(kSHC3[0], y), fs(fm(3.0f, xx), yy));
        b[10] = fm(fm(kSHC3[1], xy), z);
        b[11] = fm(fm(kSHC3[2], y), fs(fs(fm(4.0f, zz), xx), yy));
        b[12] = fm(fm(kSHC3[3], z), fs(fs(fm(2.0f, zz), fm(3.0f, xx)), fm(3.0f, yy)));
        b[13] = fm(fm(kSHC3[4], x), fs(fs(fm(4.0f, zz), xx), yy));
        b[14] = fm(fm(kSHC3[5], z), fs(xx, yy));
        b[15] = fm(fm(kSHC3[6], x), fs(xx, fm(3.0f, yy)));
      }
    }
  }
}

struct RasterArgs {
  int N, H, W, gx, gy, sh_degree, M, densify_mode, binning_mode;
  float tanfovx, tanfovy, scale_modifier;
  const float* bg;
  const float* view;
  const float* proj;
  const float* campos;
};

// ================================================================================================
// stage 1: preprocess — one thread per Gaussian
// ================================================================================================
// kSplit: the SH coefficients stay in the reference's two parameters — `shs` is _features_dc [N,1,3] and `sh_rest`
// _features_rest [N,15,3] (scene/gaussian_model.py:111-115) — instead of their concatenation.  A warp's 32 rows of
// sh_rest are one contiguous, 16-byte aligned run of 5760 bytes: the warp copies it to shared memory with coalesced
// 16-byte loads and every lane then reads its own row at a 45-word stride (odd: no bank conflicts).  The colour is
// accumulated in the same order from the same values, so it is bit-identical to the concatenated layout.
constexpr int kShRestRow = 45;                    // floats per Gaussian in _features_rest (15 coefficients x RGB)
constexpr int kShRestWarp = 32 * kShRestRow;      // floats per warp

__device__ __forceinline__ void stage_sh_rest(float* __restrict__ dst, const float* __restrict__ src, int cnt,
                                              int lane) {
  const int nf = cnt * kShRestRow;
  const float4* s4 = reinterpret_cast<const float4*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  for (int q = lane; q < (nf >> 2); q += 32) d4[q] = __ldg(s4 + q);
  for (int e = (nf & ~3) + lane; e < nf; e += 32) dst[e] = __ldg(src + e);   // (a partial last warp)
}

template <bool kSplit>
__global__ void __launch_bounds__(256)
preprocess_fwd_kernel(RasterArgs a, const float* __restrict__ means3D, const float* __restrict__ shs,
                      const float* __restrict__ sh_rest,
                      const float* __restrict__ colors_precomp, const float* __restrict__ opacities,
                      const float* __restrict__ scales, const float* __restrict__ rots,
                      const float* __restrict__ cov3D_precomp, int32_t* __restrict__ radii,
                      float2* __restrict__ means2D, float4* __restrict__ conic_opacity,
                      float4* __restrict__ rgbd, float* __restrict__ cov3D,
                      uint32_t* __restrict__ tiles_touched, uint8_t* __restrict__ clamped,
                      uint32_t* __restrict__ pair_counter) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = i < a.N;
  if (!kSplit && !valid) return;     // (kSplit: every lane stays for the warp-wide staging below)
  int radius_out = 0;
  uint32_t tiles = 0;
  float2 m2 = make_float2(0.f, 0.f);
  float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 cd = make_float4(0.f, 0.f, 0.f, 0.f);
  float cov6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  uint8_t clamp_bits = 0;
  bool visible = false;
  float p[3] = {0.f, 0.f, 0.f};
  float conx = 0.f, cony = 0.f, conz = 0.f, px = 0.f, py = 0.f, depth = 0.f;
  do {
    if (kSplit && !valid) break;
    p[0] = means3D[3 * (size_t)i]; p[1] = means3D[3 * (size_t)i + 1]; p[2] = means3D[3 * (size_t)i + 2];
    float pv[3], ph[4];
    xform4x3(a.view, p, pv);
    if (!(pv[2] > 0.2f)) break;
    xform4x4(a.proj, p, ph);
    float pw = fdv(1.0f, fa(ph[3], 0.0000001f));
    float ppx = fm(ph[0], pw), ppy = fm(ph[1], pw);
    float c6[6];
    if (cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; k++) c6[k] = cov3D_precomp[6 * (size_t)i + k];
    } else {
      float s[3] = {scales[3 * i], scales[3 * i + 1], scales[3 * i + 2]};
      float4 qv = *reinterpret_cast<const float4*>(rots + 4 * (size_t)i);
      float q[4] = {qv.x, qv.y, qv.z, qv.w};
      compute_cov3d(s, a.scale_modifier, q, c6);
    }
    Cov2D c2;
    compute_cov2d(a.W, a.H, a.tanfovx, a.tanfovy, a.view, pv, c6, c2);
    float det = ff(c2.A, c2.C, -fm(c2.B, c2.B));
    if (det == 0.0f) break;
    float det_inv = fdv(1.0f, det);
    conx = fm(c2.C, det_inv); cony = fm(-c2.B, det_inv); conz = fm(c2.A, det_inv);
    float mid = fm(0.5f, fa(c2.A, c2.C));
    float disc = cmax(0.1f, ff(mid, mid, -det));
    float sq = fsq(disc);
    float l1 = fa(mid, sq), l2 = fs(mid, sq);
    float rad_f = ceilf(fm(3.0f, fsq(cmax(l1, l2))));
    int my_radius = __float2int_rz(rad_f);
    px = ndc2pix(ppx, a.W); py = ndc2pix(ppy, a.H);
    int mnx, mny, mxx, mxy;
    get_rect(px, py, my_radius, a.gx, a.gy, mnx, mny, mxx, mxy);
    if ((mxx - mnx) * (mxy - mny) == 0) break;
#pragma unroll
    for (int k = 0; k < 6; k++) cov6[k] = c6[k];
    radius_out = my_radius;
    depth = pv[2];
    tiles = (uint32_t)((mxx - mnx) * (mxy - mny));
    visible = true;
  } while (0);
  __shared__ __align__(16) float s_rest[kSplit ? 8 : 1][kSplit ? kShRestWarp : 4];
  if (kSplit) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i0 = i - lane, cnt = min(32, a.N - i0);
    if (__any_sync(0xffffffffu, visible)) stage_sh_rest(s_rest[warp], sh_rest + (size_t)i0 * kShRestRow, cnt, lane);
    __syncwarp();
    if (!valid) return;
  }
  if (visible) {
    float rgb[3];
    if (colors_precomp) {
      rgb[0] = colors_precomp[3 * (size_t)i];
      rgb[1] = colors_precomp[3 * (size_t)i + 1];
      rgb[2] = colors_precomp[3 * (size_t)i + 2];
    } else {
      float d[3] = {fs(p[0], __ldg(a.campos)), fs(p[1], __ldg(a.campos + 1)), fs(p[2], __ldg(a.campos + 2))};
      float len = fsq(ff(d[2], d[2], ff(d[1], d[1], fm(d[0], d[0]))));
      d[0] = fdv(d[0], len); d[1] = fdv(d[1], len); d[2] = fdv(d[2], len);
      float b[16];
      sh_basis(a.sh_degree, d, b);
      int nb = (a.sh_degree + 1) * (a.sh_degree + 1);
      const float* sh = shs + (size_t)i * (kSplit ? 1 : a.M) * 3;
      float r0, r1, r2;
      if (kSplit || (a.M == 16 && (((uintptr_t)shs) & 15) == 0)) {
        float v[48];
        if (kSplit) {
          const float* row = &s_rest[threadIdx.x >> 5][(threadIdx.x & 31) * kShRestRow];
          v[0] = sh[0]; v[1] = sh[1]; v[2] = sh[2];
#pragma unroll
          for (int e = 0; e < kShRestRow; e++) v[3 + e] = (e < 3 * (nb - 1)) ? row[e] : 0.f;
        } else {
          // the usual [N,16,3] layout: 192 contiguous, 16-byte aligned bytes per Gaussian -> up to twelve 16-byte loads
          // issued together instead of 48 dependent scalar loads; same FMA order, so the colours stay bit-identical
          const float4* s4 = reinterpret_cast<const float4*>(sh);
#pragma unroll
          for (int q = 0; q < 12; q++) {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (4 * q < 3 * nb) t = s4[q];
            v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
          }
        }
        r0 = fm(b[0], v[0]); r1 = fm(b[0], v[1]); r2 = fm(b[0], v[2]);
#pragma unroll
        for (int k = 1; k < 16; k++) {
          if (k < nb) {
            r0 = ff(b[k], v[3 * k], r0);
            r1 = ff(b[k], v[3 * k + 1], r1);
            r2 = ff(b[k], v[3 * k + 2], r2);
          }
        }
      } else {
        r0 = fm(b[0], sh[0]); r1 = fm(b[0], sh[1]); r2 = fm(b[0], sh[2]);
        for (int k = 1; k < nb; k++) {
          r0 = ff(b[k], sh[3 * k], r0);
          r1 = ff(b[k], sh[3 * k + 1], r1);
          r2 = ff(b[k], sh[3 * k + 2], r2);
        }
      }
      r0 = fa(r0, 0.5f); r1 = fa(r1, 0.5f); r2 = fa(r2, 0.5f);
      clamp_bits = (r0 < 0.0f ? 1 : 0) | (r1 < 0.0f ? 2 : 0) | (r2 < 0.0f ? 4 : 0);
      rgb[0] = cmax(r0, 0.0f); rgb[1] = cmax(r1, 0.0f); rgb[2] = cmax(r2, 0.0f);
    }
    m2 = make_float2(px, py);
    co = make_float4(conx, cony, conz, opacities[i]);
    cd = make_float4(rgb[0], rgb[1], rgb[2], depth);
  }
  if (!visible) clamp_bits = 0;
  radii[i] = radius_out;
  means2D[i] = m2;
  conic_opacity[i] = co;
  rgbd[i] = cd;
#pragma unroll
  for (int k = 0; k < 6; k++) cov3D[6 * (size_t)i + k] = cov6[k];
  tiles_touched[i] = tiles;
  clamped[i] = clamp_bits;
  if (pair_counter) {   // tile binning: R = sum of tiles_touched without a scan over the Gaussians
    const unsigned m = __activemask();
    const uint32_t sum = __reduce_add_sync(m, tiles);
    if ((threadIdx.x & 31) == (__ffs(m) - 1) && sum) atomicAdd(pair_counter, sum);
  }
}

__global__ void finish_scan_kernel(int N, const uint32_t* __restrict__ offsets, uint32_t* header) {
  header[0] = N > 0 ? offsets[N - 1] : 0;
  header[1] = 0;
}

// ================================================================================================
// stage 2: binning (App. A.2)
// ================================================================================================
__global__ void __launch_bounds__(256)
duplicate_with_keys_kernel(int N, int gx, int gy, uint64_t cap, const float2* __restrict__ means2D,
                           const float4* __restrict__ rgbd, const uint32_t* __restrict__ offsets,
                           const int32_t* __restrict__ radii, uint64_t* __restrict__ keys,
                           uint32_t* __restrict__ vals, uint32_t* header) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  int r = radii[i];
  if (r <= 0) return;
  uint32_t off = i ? offsets[i - 1] : 0u;
  uint32_t end = offsets[i];
  if ((uint64_t)end > cap) { header[1] = 1; return; }
  float2 m = means2D[i];
  int mnx, mny, mxx, mxy;
  get_rect(m.x, m.y, r, gx, gy, mnx, mny, mxx, mxy);
  uint32_t db = __float_as_uint(rgbd[i].w);
  for (int y = mny; y < mxy; y++)
    for (int x = mnx; x < mxx; x++) {
      uint64_t key = ((uint64_t)(uint32_t)(y * gx + x) << 32) | db;
      keys[off] = key;
      vals[off] = (uint32_t)i;
      off++;
    }
}

// keys in [R,cap) are set to all-ones so they sort behind every real key
__global__ void pad_keys_kernel(uint64_t cap, const uint32_t* __restrict__ header,
                                uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t R = header[0];
  if (j >= cap || j < R) return;
  keys[j] = ~0ull;
  vals[j] = 0u;
}

__global__ void __launch_bounds__(256)
identify_tile_ranges_kernel(uint64_t cap, const uint32_t* __restrict__ header,
                            const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
  uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t R = header[0];
  if (R > cap) R = 0;
  if (j >= R) return;
  uint32_t tile = (uint32_t)(keys[j] >> 32);
  if (j == 0) ranges[tile].x = 0;
  else {
    uint32_t prev = (uint32_t)(keys[j - 1] >> 32);
    if (tile != prev) { ranges[prev].y = (uint32_t)j; ranges[tile].x = (uint32_t)j; }
  }
  if (j == R - 1) ranges[tile].y = (uint32_t)R;
}

// ------------------------------------------------------------------------------------------------
// Tile binning (binning_mode 0, the default): the same sorted (tile | depth) list, point list and tile ranges as
// duplicateWithKeys + 64-bit radix sort + identifyTileRanges, without sorting the tile bits and without six
// 8-bit passes over all pairs:
//   count   : one atomic per (tile, Gaussian) pair on the tile's counter
//   scan    : exclusive scan of the tile counters = the tile ranges (empty tiles stay (0,0) as in the reference)
//   scatter : each pair takes the next slot of its tile's segment and stores (depth bits << 32 | Gaussian index)
//   sort    : one CTA per tile sorts its segment in shared memory by that 64-bit key.  The keys of a tile are
//             unique, and the reference's stable sort of (tile | depth) leaves equal keys in emission order =
//             ascending Gaussian index, so the result is bit-identical to the radix path.
// Segments longer than kTileSortCap are sorted in chunks and merged through global memory by the same CTA.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_count_kernel(int N, int gx, int gy, const float2* __restrict__ means2D, const int32_t* __restrict__ radii,
                  uint32_t* __restrict__ tile_count) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  int r = radii[i];
  if (r <= 0) return;
  float2 m = means2D[i];
  int mnx, mny, mxx, mxy;
  get_rect(m.x, m.y, r, gx, gy, mnx, mny, mxx, mxy);
  for (int y = mny; y < mxy; y++)
    for (int x = mnx; x < mxx; x++) atomicAdd(&tile_count[(size_t)(y * gx + x) * kTileCtrStride], 1u);
}

// single CTA: exclusive scan of the tile counters -> ranges, cursors; checks R against the capacity
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int tiles, uint64_t cap, const uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_cursor,
                 uint2* __restrict__ ranges, uint32_t* __restrict__ header) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  const int per = (tiles + 1023) / 1024;
  const int t0 = tid * per, t1 = min(tiles, t0 + per);
  uint32_t local = 0;
  for (int t = t0; t < t1; t++) local += tile_count[(size_t)t * kTileCtrStride];
  uint32_t inc = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = s_warp[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += v; }
    s_warp[lane] = winc - w;
    if (lane == 31) s_carry = winc;
  }
  __syncthreads();
  uint32_t off = s_warp[warp] + inc - local;
  const uint32_t R = s_carry;
  const bool overflow = (uint64_t)R > cap;
  for (int t = t0; t < t1; t++) {
    const uint32_t c = tile_count[(size_t)t * kTileCtrStride];
    tile_cursor[(size_t)t * kTileCtrStride] = off;
    ranges[t] = (c && !overflow) ? make_uint2(off, off + c) : make_uint2(0u, 0u);
    off += c;
  }
  if (tid == 0) { header[0] = R; header[1] = overflow ? 1u : 0u; }
}

// The slot of a pair is the return value of an atomic.  One thread per Gaussian walking its own tiles makes a latency
// chain as long as the largest splat of the warp (measured 105 us at C2, 62 us with eight atomics in flight per
// thread).  Here a warp flattens the pairs of its 32 Gaussians: pair p of the warp goes to lane p % 32, which finds
// the owning Gaussian in the warp's prefix sums by shuffles; every lane keeps four independent atomics in flight.
__global__ void __launch_bounds__(256)
tile_scatter_kernel(int N, int gx, int gy, const float2* __restrict__ means2D, const float4* __restrict__ rgbd,
                    const int32_t* __restrict__ radii, uint32_t* __restrict__ tile_cursor,
                    uint64_t* __restrict__ seg, const uint32_t* __restrict__ header) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  if (header[1]) return;
  int mnx = 0, mny = 0, w = 0, cnt = 0;
  uint32_t depth_bits = 0;
  if (i < N) {
    const int r = radii[i];
    if (r > 0) {
      const float2 m = means2D[i];
      int mxx, mxy;
      get_rect(m.x, m.y, r, gx, gy, mnx, mny, mxx, mxy);
      w = mxx - mnx;
      cnt = w * (mxy - mny);
      depth_bits = __float_as_uint(rgbd[i].w);
    }
  }
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  for (int p0 = 0; p0 < total; p0 += 128) {
    uint32_t pos[4], owner[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int p = p0 + k * 32 + lane;
      // owner = first lane whose inclusive prefix exceeds p (lower bound over 32 lanes, 5 shuffle steps)
      int lo = 0;
#pragma unroll
      for (int step = 16; step > 0; step >>= 1) {
        const int v = __shfl_sync(0xffffffffu, incl, lo + step - 1);
        if (v <= p) lo += step;
      }
      const int o_incl = __shfl_sync(0xffffffffu, incl, lo), o_cnt = __shfl_sync(0xffffffffu, cnt, lo);
      const int o_w = __shfl_sync(0xffffffffu, w, lo), o_x = __shfl_sync(0xffffffffu, mnx, lo);
      const int o_y = __shfl_sync(0xffffffffu, mny, lo);
      owner[k] = (uint32_t)lo;
      pos[k] = 0xFFFFFFFFu;
      if (p < total) {
        const int t = p - (o_incl - o_cnt);
        const int y = o_y + t / o_w, x = o_x + t % o_w;
        pos[k] = atomicAdd(&tile_cursor[(size_t)(y * gx + x) * kTileCtrStride], 1u);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t d = __shfl_sync(0xffffffffu, depth_bits, owner[k]);
      if (pos[k] != 0xFFFFFFFFu)
        seg[pos[k]] = ((uint64_t)d << 32) | (uint32_t)(i - lane + (int)owner[k]);
    }
  }
}

static constexpr int kTileSortCap = 4096;   // keys sorted in shared memory at once (32 KB)

// bitonic sort of n_pad (power of two <= kTileSortCap) 64-bit keys in shared memory, ascending
// (steps with partner distance < 64 stay inside an aligned block of 64 keys: one warp runs all of them for its
// blocks back to back with __syncwarp only; block-wide barriers are needed for the longer-distance steps alone)
__device__ __forceinline__ void smem_bitonic(uint64_t* k, int n_pad) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int size = 2; size <= n_pad; size <<= 1) {
    int j = size >> 1;
    for (; j >= 64; j >>= 1) {
      for (int t = threadIdx.x; t < (n_pad >> 1); t += blockDim.x) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), p = i | j;
        const uint64_t a = k[i], b = k[p];
        if ((a > b) == ((i & size) == 0)) { k[i] = b; k[p] = a; }
      }
      __syncthreads();
    }
    for (int blk = warp; blk < ((n_pad + 63) >> 6); blk += nwarps) {
      const int base = blk << 6;
      for (int jj = j; jj > 0; jj >>= 1) {
        const int i = base + (((lane & ~(jj - 1)) << 1) | (lane & (jj - 1))), p = i | jj;
        if (p < n_pad) {
          const uint64_t a = k[i], b = k[p];
          if ((a > b) == ((i & size) == 0)) { k[i] = b; k[p] = a; }
        }
        __syncwarp();
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
tile_sort_kernel(const uint2* __restrict__ ranges, uint64_t* seg, uint64_t* tmp, uint64_t* keys_sorted,
                 uint32_t* __restrict__ point_list) {
  extern __shared__ __align__(16) uint64_t s_keys[];
  const uint32_t tile = blockIdx.x;
  const uint2 rg = ranges[tile];
  const int n = (int)(rg.y - rg.x);
  if (n <= 0) return;
  uint64_t* src = seg + rg.x;
  if (n <= kTileSortCap) {
    int n_pad = 2;
    while (n_pad < n) n_pad <<= 1;
    for (int i = threadIdx.x; i < n_pad; i += blockDim.x) s_keys[i] = i < n ? src[i] : ~0ull;
    __syncthreads();
    smem_bitonic(s_keys, n_pad);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const uint64_t r = s_keys[i];
      keys_sorted[rg.x + i] = ((uint64_t)tile << 32) | (r >> 32);
      point_list[rg.x + i] = (uint32_t)r;
    }
    return;
  }
  // long segment: sort chunks of kTileSortCap in shared memory, then merge runs pairwise through global memory
  uint64_t* dst = tmp + rg.x;
  for (int c0 = 0; c0 < n; c0 += kTileSortCap) {
    const int m = min(kTileSortCap, n - c0);
    for (int i = threadIdx.x; i < kTileSortCap; i += blockDim.x) s_keys[i] = i < m ? src[c0 + i] : ~0ull;
    __syncthreads();
    smem_bitonic(s_keys, kTileSortCap);
    for (int i = threadIdx.x; i < m; i += blockDim.x) src[c0 + i] = s_keys[i];
    __syncthreads();
  }
  for (int run = kTileSortCap; run < n; run <<= 1) {
    __threadfence_block();
    __syncthreads();
    // each thread produces a contiguous slice of every merged pair of runs (merge path by binary search)
    for (int base = 0; base < n; base += 2 * run) {
      const int la = min(run, n - base), lb = max(0, min(run, n - base - run));
      const uint64_t* A = src + base;
      const uint64_t* B = src + base + la;
      const int total = la + lb;
      const int per = (total + (int)blockDim.x - 1) / (int)blockDim.x;
      const int o0 = min(total, (int)threadIdx.x * per), o1 = min(total, o0 + per);
      if (o0 < o1) {
        int lo = max(0, o0 - lb), hi = min(o0, la);   // elements taken from A among the first o0 outputs
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (A[mid] <= B[o0 - mid - 1]) lo = mid + 1; else hi = mid;
        }
        int ia = lo, ib = o0 - lo;
        for (int o = o0; o < o1; o++) {
          const bool takeA = ib >= lb || (ia < la && A[ia] <= B[ib]);
          dst[base + o] = takeA ? A[ia++] : B[ib++];
        }
      }
    }
    __threadfence_block();
    __syncthreads();
    uint64_t* t = src; src = dst; dst = t;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const uint64_t r = src[i];
    keys_sorted[rg.x + i] = ((uint64_t)tile << 32) | (r >> 32);
    point_list[rg.x + i] = (uint32_t)r;
  }
}

// Conservative early-out for the blend loops.  A Gaussian contributes to a pixel only if
// alpha = opacity * exp(power) >= 1/255.  For power < -log(255 * opacity) - 1e-4 the exact product is below
// (1 - 9e-5)/255 while det_exp, logf and the multiplications are each accurate to ~1e-6 relative, so the
// exact test `alpha < 1/255` is guaranteed to reject as well: results stay bit-identical to the oracle.
// opacity <= 0 can never reach 1/255 (threshold +inf); a NaN opacity yields a NaN threshold = never early-out.
__device__ __forceinline__ float reject_threshold(float opacity) {
  if (opacity <= 0.0f) return __int_as_float(0x7f800000);
  return -logf(255.0f * opacity) - 1e-4f;
}

// Which of the tile's eight warps (8x4 pixel blocks; warp w covers x-half w&1, y-quarter w>>1) can a Gaussian
// reach?  {power >= thr} is the ellipse a*dx^2 + 2b*dx*dy + c*dy^2 <= -2*thr, whose bounding box has half
// extents sqrt(-2*thr*c/det), sqrt(-2*thr*a/det).  The box is widened (0.1 % + 0.01 px) far beyond the fp32
// rounding of `power`, so a warp whose bit is clear would have rejected the Gaussian at every pixel with the
// exact per-pixel tests: skipping it leaves the results bit-identical.  Anything non-finite keeps all warps.
__device__ __forceinline__ uint32_t warp_reach_mask(float mx, float my, float a, float b, float c, float thr,
                                                    float tile_x0, float tile_y0) {
  const float det = a * c - b * b;
  const float inv = (-2.0f * thr) / det;
  const float hx = sqrtf(c * inv) * 1.001f + 0.01f, hy = sqrtf(a * inv) * 1.001f + 0.01f;
  if (!(det > 0.0f) || !(hx < 1e30f) || !(hy < 1e30f)) return 0xFFu;
  const float x0 = mx - hx - tile_x0, x1 = mx + hx - tile_x0;
  const float y0 = my - hy - tile_y0, y1 = my + hy - tile_y0;
  const uint32_t xb = ((x1 >= 0.0f && x0 <= 7.0f) ? 1u : 0u) | ((x1 >= 8.0f && x0 <= 15.0f) ? 2u : 0u);
  uint32_t m = 0;
#pragma unroll
  for (int q = 0; q < 4; q++)
    if (y1 >= (float)(4 * q) && y0 <= (float)(4 * q + 3)) m |= xb << (2 * q);
  return m;
}
// Transpose the per-Gaussian reach masks of a staged batch into one 256-bit bitmap per pixel warp.
__device__ __forceinline__ void publish_reach(uint32_t m, uint32_t (*s_bits)[8]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int p = 0; p < 8; p++) {
    const uint32_t word = __ballot_sync(0xffffffffu, (m >> p) & 1u);
    if (lane == p) s_bits[p][warp] = word;
  }
}

// ================================================================================================
// stage 3: render forward (App. A.3) — one CTA per 16x16 tile, one thread per pixel
// ================================================================================================
__global__ void __launch_bounds__(kBlock)
render_fwd_kernel(int H, int W, int gx, const float* __restrict__ bg,
                  const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                  const float4* __restrict__ rgbd, float* __restrict__ out_color,
                  float* __restrict__ out_depth, float* __restrict__ final_T,
                  uint32_t* __restrict__ n_contrib) {
  __shared__ float4 s_xy[kBlock];   // (x, y, reject threshold on power, -)
  __shared__ float4 s_co[kBlock];
  __shared__ float4 s_cd[kBlock];
  __shared__ uint32_t s_bits[8][8];   // [pixel warp][staging warp]: Gaussians of the batch that can reach the warp
  int tile = blockIdx.y * gx + blockIdx.x;
  // a warp covers an 8x4 pixel block (compact footprint: fewer warps touched per splat than 16x2 strips)
  int lx = (threadIdx.x & 7) | ((threadIdx.x >> 2) & 8), ly = ((threadIdx.x >> 3) & 3) | ((threadIdx.x >> 4) & 12);
  int pxi = blockIdx.x * kTile + lx, pyi = blockIdx.y * kTile + ly;
  bool inside = pxi < W && pyi < H;
  float pfx = (float)pxi, pfy = (float)pyi;
  uint2 range = ranges[tile];
  int todo = (int)(range.y - range.x);
  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t contributor = 0, last = 0;
  for (int base = 0; base < todo; base += kBlock) {
    if (__syncthreads_count(done) == kBlock) break;
    int idx = base + threadIdx.x;
    uint32_t reach = 0;
    if (idx < todo) {
      uint32_t id = point_list[range.x + idx];
      float2 m = means2D[id];
      float4 co = conic_opacity[id];
      const float thr = reject_threshold(co.w);
      s_xy[threadIdx.x] = make_float4(m.x, m.y, thr, 0.f);
      s_co[threadIdx.x] = co;
      s_cd[threadIdx.x] = rgbd[id];
      reach = warp_reach_mask(m.x, m.y, co.x, co.y, co.z, thr, (float)(blockIdx.x * kTile), (float)(blockIdx.y * kTile));
    }
    publish_reach(reach, s_bits);
    __syncthreads();
    for (int wi = 0; wi < kBlock / 32; wi++) {
      uint32_t word = s_bits[threadIdx.x >> 5][wi];
      if (word == 0u) continue;
      if (__all_sync(0xffffffffu, done)) break;
      while (word) {
      const int j = wi * 32 + __ffs(word) - 1;
      word &= word - 1u;
      if (done) continue;
      contributor = (uint32_t)(base + j + 1);
      float4 xy = s_xy[j];
      float4 co = s_co[j];
      float dx = fs(xy.x, pfx), dy = fs(xy.y, pfy);
      float q = ff(fm(co.z, dy), dy, fm(fm(co.x, dx), dx));
      float power = ff(-0.5f, q, -fm(fm(co.y, dx), dy));
      // power < threshold: alpha is certainly below 1/255 (see reject_threshold), so the exact test below
      // would skip this Gaussian too — decided without evaluating the exponential
      if (power > 0.0f || power < xy.z) continue;
      float alpha = cmin(0.99f, fm(co.w, det_exp(power)));
      if (alpha < (float)(1.0 / 255.0)) continue;
      float test_T = fm(T, fs(1.0f, alpha));
      if (test_T < 0.0001f) { done = true; continue; }
      float4 cd = s_cd[j];
      float w = fm(alpha, T);
      C0 = ff(cd.x, w, C0);
      C1 = ff(cd.y, w, C1);
      C2 = ff(cd.z, w, C2);
      D = ff(cd.w, w, D);
      T = test_T;
      last = contributor;
      }
    }
  }
  if (inside) {
    size_t pix = (size_t)pyi * W + pxi;
    size_t P = (size_t)H * W;
    final_T[pix] = T;
    n_contrib[pix] = last;
    out_color[pix] = ff(T, __ldg(bg), C0);
    out_color[P + pix] = ff(T, __ldg(bg + 1), C1);
    out_color[2 * P + pix] = ff(T, __ldg(bg + 2), C2);
    out_depth[pix] = D;
  }
}

// ================================================================================================
// stage 4: render backward (App. A.4).  The twelve per-(pixel,Gaussian) gradient terms are reduced over
// the 32 pixels of a warp with a 13-shuffle split butterfly (each step halves the number of live values
// per lane instead of reducing every value five times), added into a per-batch shared-memory accumulator
// shared by the 8 warps of the tile, and flushed once per (tile,Gaussian) with three 16-byte vector
// reductions (RED.128) — instead of the ~10 scalar global atomics per (pixel,Gaussian) of the classic
// kernel.  Backward is held to 1e-3 relative, not bit-exactness, so it uses MUFU ex2/rcp.
// ================================================================================================
__device__ __forceinline__ void red_add_v4(float4* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ float fast_rcp(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float fast_ex2(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// one butterfly step: lanes with `up` keep hi and send lo, the others keep lo and send hi
__device__ __forceinline__ float bfly(float lo, float hi, bool up, int mask) {
  float keep = up ? hi : lo, send = up ? lo : hi;
  return keep + __shfl_xor_sync(0xffffffffu, send, mask);
}

static constexpr int kAccStride = 12;  // floats per Gaussian in the shared accumulator

// kDepth: a depth cotangent exists (the reference never back-propagates through depth, so the usual instantiation
// drops the fourth channel's recurrence and its gradient term); kDensify: semantics of the densify buffer.
template <bool kDepth, int kDensify>
__global__ void __launch_bounds__(kBlock)
render_bwd_kernel(int H, int W, int gx, const float* __restrict__ bg,
                  const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                  const float4* __restrict__ rgbd, const float* __restrict__ final_T,
                  const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                  const float* __restrict__ dL_ddepth, float4* __restrict__ acc_mean,
                  float4* __restrict__ acc_conic, float4* __restrict__ acc_color) {
  __shared__ uint32_t s_id[kBlock];
  __shared__ float2 s_xy[kBlock];
  __shared__ float4 s_co[kBlock];   // conic scaled by log2(e), opacity
  __shared__ float4 s_cd[kBlock];
  __shared__ __align__(16) float s_acc[kBlock * kAccStride];
  __shared__ int s_max[kBlock / 32];
  __shared__ uint32_t s_bits[8][8];
  const int tile = blockIdx.y * gx + blockIdx.x;
  const int lane = threadIdx.x & 31;
  const int lx = (threadIdx.x & 7) | ((threadIdx.x >> 2) & 8), ly = ((threadIdx.x >> 3) & 3) | ((threadIdx.x >> 4) & 12);
  const int pxi = blockIdx.x * kTile + lx, pyi = blockIdx.y * kTile + ly;
  const bool inside = pxi < W && pyi < H;
  const float pfx = (float)pxi, pfy = (float)pyi;
  const uint2 range = ranges[tile];
  int todo = (int)(range.y - range.x);
  const size_t P = (size_t)H * W;
  const size_t pix = inside ? (size_t)pyi * W + pxi : 0;
  const float T_final = inside ? final_T[pix] : 0.f;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix] : 0;
  float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f, dLd = 0.f;
  if (inside) {
    dLp0 = dL_dpix[pix]; dLp1 = dL_dpix[P + pix]; dLp2 = dL_dpix[2 * P + pix];
    if (kDepth) dLd = dL_ddepth[pix];
  }
  const float bgT = -T_final * (__ldg(bg) * dLp0 + __ldg(bg + 1) * dLp1 + __ldg(bg + 2) * dLp2);
  float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, ard = 0.f;
  float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, ld = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * (float)W, ddely_dy = 0.5f * (float)H;
  const float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
  // butterfly bookkeeping: which of the 12 sums this lane ends up holding
  const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4, up2 = lane & 2;
  const int my_slot = (up16 ? 6 : 0) + (up8 ? 3 : 0) + (up4 ? 2 : (up2 ? 1 : 0));
  const bool slot_owner = !(lane & 1) && !(up4 && up2);
  // the block only needs the first max(last_contributor) entries of the list; each warp additionally drops the
  // entries past its own maximum by clearing their bits in its reach bitmap (free, unlike a per-iteration test)
  const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
  {
    int m = warp_last;
    if (lane == 0) s_max[threadIdx.x >> 5] = m;
    __syncthreads();
    int block_last = 0;
#pragma unroll
    for (int w = 0; w < kBlock / 32; w++) block_last = max(block_last, s_max[w]);
    todo = min(todo, block_last);
  }
  for (int base = 0; base < todo; base += kBlock) {
    __syncthreads();
    const int idx = base + threadIdx.x;
    uint32_t reach = 0;
    if (idx < todo) {
      uint32_t id = point_list[range.x + (todo - 1 - idx)];
      float4 co = conic_opacity[id];
      const float2 m = means2D[id];
      s_id[threadIdx.x] = id;
      s_xy[threadIdx.x] = m;
      s_co[threadIdx.x] = make_float4(co.x * kLog2e, co.y * kLog2e, co.z * kLog2e, co.w);
      s_cd[threadIdx.x] = rgbd[id];
      reach = warp_reach_mask(m.x, m.y, co.x, co.y, co.z, reject_threshold(co.w), (float)(blockIdx.x * kTile),
                              (float)(blockIdx.y * kTile));
    }
    publish_reach(reach, s_bits);
    {
      float4* z = (float4*)(s_acc + threadIdx.x * kAccStride);
      z[0] = make_float4(0.f, 0.f, 0.f, 0.f); z[1] = z[0]; z[2] = z[0];
    }
    __syncthreads();
    const int jmin = todo - warp_last - base;   // entries j < jmin have contributor >= warp_last: invalid for every lane
    for (int wi = 0; wi < kBlock / 32; wi++) {
      uint32_t word = s_bits[threadIdx.x >> 5][wi];
      const int lo = jmin - wi * 32;
      if (lo >= 32) continue;
      if (lo > 0) word &= 0xffffffffu << lo;
      while (word) {
      const int j = wi * 32 + __ffs(word) - 1;
      word &= word - 1u;
      const int contributor = todo - 1 - (base + j);
      const float2 xy = s_xy[j];
      const float4 co = s_co[j];
      const float dx = xy.x - pfx, dy = xy.y - pfy;
      const float cdx = co.x * dx, cdy = co.z * dy;
      const float p2 = -0.5f * (cdx * dx + cdy * dy) - co.y * dx * dy;  // power * log2(e)
      // (a conservative threshold test ahead of the exponential, as in the forward kernel, was measured slower
      // here: MUFU.EX2 is two instructions, the extra ballot costs more)
      const float G = fast_ex2(fminf(p2, 0.0f));
      const float alpha = fminf(0.99f, co.w * G);
      const bool valid = contributor < last_contributor && !(p2 > 0.0f) && !(alpha < (float)(1.0 / 255.0));
      if (__ballot_sync(0xffffffffu, valid) == 0u) continue;
      float v[12];
#pragma unroll
      for (int k = 0; k < 12; k++) v[k] = 0.f;
      if (valid) {
        const float4 cd = s_cd[j];
        const float inv = fast_rcp(1.0f - alpha);
        T *= inv;
        const float dch = alpha * T;
        ar0 = last_alpha * lc0 + (1.0f - last_alpha) * ar0; lc0 = cd.x;
        ar1 = last_alpha * lc1 + (1.0f - last_alpha) * ar1; lc1 = cd.y;
        ar2 = last_alpha * lc2 + (1.0f - last_alpha) * ar2; lc2 = cd.z;
        float dL_dalpha = (cd.x - ar0) * dLp0 + (cd.y - ar1) * dLp1 + (cd.z - ar2) * dLp2;
        if (kDepth) {
          ard = last_alpha * ld + (1.0f - last_alpha) * ard;  ld = cd.w;
          dL_dalpha += (cd.w - ard) * dLd;
        }
        dL_dalpha = dL_dalpha * T + bgT * inv;
        last_alpha = alpha;
        const float dL_dG = co.w * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        // s_co holds conic * log2(e): undo the scale once on the (conic . d) terms
        const float cy_dx = co.y * dx, cy_dy = co.y * dy;
        const float dG_ddelx = -kLn2 * (G * cdx + G * cy_dy);
        const float dG_ddely = -kLn2 * (G * cdy + G * cy_dx);
        const float g_mx = dL_dG * dG_ddelx * ddelx_dx;
        const float g_my = dL_dG * dG_ddely * ddely_dy;
        v[0] = g_mx; v[1] = g_my;
        if (kDensify == 1) { v[2] = g_mx; v[3] = g_my; }
        else if (kDensify == 2) { v[2] = fabsf(g_mx); v[3] = fabsf(g_my); }
        else { v[2] = sqrtf(g_mx * g_mx + g_my * g_my); }
        const float hG = -0.5f * dL_dG;
        v[4] = hG * gdx * dx;
        v[5] = hG * gdx * dy;
        v[6] = hG * gdy * dy;
        v[7] = G * dL_dalpha;
        v[8] = dch * dLp0; v[9] = dch * dLp1; v[10] = dch * dLp2; v[11] = kDepth ? dch * dLd : 0.f;
      }
      // split butterfly: 12 -> 6 -> 3 -> 2 -> 1 values per lane, 13 shuffles
      float a6[6], b3[3];
#pragma unroll
      for (int k = 0; k < 6; k++) a6[k] = bfly(v[k], v[k + 6], up16, 16);
#pragma unroll
      for (int k = 0; k < 3; k++) b3[k] = bfly(a6[k], a6[k + 3], up8, 8);
      const float c0 = bfly(b3[0], b3[2], up4, 4);
      const float c1 = bfly(b3[1], 0.f, up4, 4);
      float d = bfly(c0, c1, up2, 2);
      d += __shfl_xor_sync(0xffffffffu, d, 1);
      if (slot_owner) atomicAdd(s_acc + j * kAccStride + my_slot, d);
      }
    }
    __syncthreads();
    if (idx < todo) {
      const float4* r = (const float4*)(s_acc + threadIdx.x * kAccStride);
      const float4 m = r[0], c = r[1], k = r[2];
      const bool nz = (m.x != 0.f) | (m.y != 0.f) | (m.z != 0.f) | (m.w != 0.f) | (c.x != 0.f) | (c.y != 0.f) |
                      (c.z != 0.f) | (c.w != 0.f) | (k.x != 0.f) | (k.y != 0.f) | (k.z != 0.f) | (k.w != 0.f);
      if (nz) {
        const uint32_t id = s_id[threadIdx.x];
        red_add_v4(acc_mean + id, m.x, m.y, m.z, m.w);
        red_add_v4(acc_conic + id, c.x, c.y, c.z, c.w);
        red_add_v4(acc_color + id, k.x, k.y, k.z, k.w);
      }
    }
  }
}

// ================================================================================================
// stage 5: preprocess backward (cov2D bwd + mean/colour/cov3D chain) — one thread per Gaussian
// ================================================================================================
// kSplit (see preprocess_fwd_kernel): `shs` / `sh_rest` are _features_dc / _features_rest, and dL/dSH is ADDED to
// `dL_dsh` ([N,1,3], the .grad of _features_dc) and `dL_dsh_rest` ([N,15,3], the .grad of _features_rest): the warp
// stages its 32 gradient rows in the shared-memory block the coefficients were read from and adds the contiguous
// 5760-byte run to global memory with coalesced 16-byte read-add-writes.
template <bool kSplit>
__global__ void __launch_bounds__(256)
preprocess_bwd_kernel(RasterArgs a, const float* __restrict__ means3D, const float* __restrict__ shs,
                      const float* __restrict__ sh_rest,
                      const float* __restrict__ colors_precomp, const float* __restrict__ scales,
                      const float* __restrict__ rots, const int32_t* __restrict__ radii,
                      const float* __restrict__ cov3D, const uint8_t* __restrict__ clamped,
                      const float4* __restrict__ acc_mean, const float4* __restrict__ acc_conic,
                      const float4* __restrict__ acc_color, float* __restrict__ dL_dmeans3D,
                      float* __restrict__ dL_dmeans2D, float* __restrict__ dL_dmeans2D_densify,
                      float* __restrict__ dL_dsh, float* __restrict__ dL_dsh_rest,
                      float* __restrict__ dL_dcolors_precomp,
                      float* __restrict__ dL_dopacities, float* __restrict__ dL_dscales,
                      float* __restrict__ dL_drots, float* __restrict__ dL_dcov3D) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = i < a.N;
  if (!kSplit && !valid) return;
  bool vis = valid && radii[i] > 0;
  __shared__ __align__(16) float s_rest[kSplit ? 8 : 1][kSplit ? kShRestWarp : 4];
  if (kSplit) {   // every lane of the warp takes part in the staging, then the lanes past N leave
    const int lane0 = threadIdx.x & 31;
    const int i0 = i - lane0, cnt0 = min(32, a.N - i0);
    if (__any_sync(0xffffffffu, vis)) stage_sh_rest(s_rest[threadIdx.x >> 5], sh_rest + (size_t)i0 * kShRestRow, cnt0, lane0);
    __syncwarp();
    if (!valid) return;
  }
  size_t si = (size_t)i;
  float4 gm = vis ? acc_mean[i] : make_float4(0, 0, 0, 0);
  float4 gc = vis ? acc_conic[i] : make_float4(0, 0, 0, 0);
  float4 gcol = vis ? acc_color[i] : make_float4(0, 0, 0, 0);
  if (dL_dmeans2D) { dL_dmeans2D[3 * si] = gm.x; dL_dmeans2D[3 * si + 1] = gm.y; dL_dmeans2D[3 * si + 2] = 0.f; }
  if (dL_dmeans2D_densify) {
    dL_dmeans2D_densify[3 * si] = gm.z; dL_dmeans2D_densify[3 * si + 1] = gm.w; dL_dmeans2D_densify[3 * si + 2] = 0.f;
  }
  if (dL_dopacities) dL_dopacities[si] = gc.w;
  float dmean[3] = {0.f, 0.f, 0.f};
  float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dsc[3] = {0.f, 0.f, 0.f};
  float dq[4] = {0.f, 0.f, 0.f, 0.f};
  int nb = (a.sh_degree + 1) * (a.sh_degree + 1);
  // dL/dSH of a Gaussian is basis[k] * dRGB[c]: kept in registers and written after the branch by the whole warp
  // (invisible Gaussians write zeros), through shared memory so that the stores are contiguous runs instead of
  // 48 scalar stores at a 192-byte lane stride
  float shb[16], shg[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 16; k++) shb[k] = 0.f;
  if (!vis) {
    if (dL_dcolors_precomp) for (int k = 0; k < 3; k++) dL_dcolors_precomp[3 * si + k] = 0.f;
  } else {
    float p[3] = {means3D[3 * si], means3D[3 * si + 1], means3D[3 * si + 2]};
    // ---- cov2D backward ----
    float pv[3];
    xform4x3(a.view, p, pv);
    float c6[6];
#pragma unroll
    for (int k = 0; k < 6; k++) c6[k] = cov3D[6 * si + k];
    Cov2D c2;
    compute_cov2d(a.W, a.H, a.tanfovx, a.tanfovy, a.view, pv, c6, c2);
    float xg = c2.clampx ? 0.f : 1.f, yg = c2.clampy ? 0.f : 1.f;
    float A = c2.A, B = c2.B, C = c2.C;
    const float* T = c2.T;
    float denom = A * C - B * B;
    float d2 = 1.0f / (denom * denom + 0.0000001f);
    float dT[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (d2 != 0.0f) {
      float dL_da = d2 * (-C * C * gc.x + 2.0f * B * C * gc.y + (denom - A * C) * gc.z);
      float dL_dc = d2 * (-A * A * gc.z + 2.0f * A * B * gc.y + (denom - A * C) * gc.x);
      float dL_db = d2 * 2.0f * (B * C * gc.x - (denom + 2.0f * B * B) * gc.y + A * B * gc.z);
      dcov[0] = T[0] * T[0] * dL_da + T[3] * T[3] * dL_dc + T[0] * T[3] * dL_db;
      dcov[3] = T[1] * T[1] * dL_da + T[4] * T[4] * dL_dc + T[1] * T[4] * dL_db;
      dcov[5] = T[2] * T[2] * dL_da + T[5] * T[5] * dL_dc + T[2] * T[5] * dL_db;
      dcov[1] = 2.0f * (T[0] * T[1] * dL_da + T[3] * T[4] * dL_dc) + (T[0] * T[4] + T[1] * T[3]) * dL_db;
      dcov[2] = 2.0f * (T[0] * T[2] * dL_da + T[3] * T[5] * dL_dc) + (T[0] * T[5] + T[2] * T[3]) * dL_db;
      dcov[4] = 2.0f * (T[1] * T[2] * dL_da + T[4] * T[5] * dL_dc) + (T[1] * T[5] + T[2] * T[4]) * dL_db;
      float S[9] = {c6[0], c6[1], c6[2], c6[1], c6[3], c6[4], c6[2], c6[4], c6[5]};
#pragma unroll
      for (int k = 0; k < 3; k++) {
        float s0 = S[3 * k] * T[0] + S[3 * k + 1] * T[1] + S[3 * k + 2] * T[2];
        float s1 = S[3 * k] * T[3] + S[3 * k + 1] * T[4] + S[3 * k + 2] * T[5];
        dT[k] = 2.0f * s0 * dL_da + s1 * dL_db;
        dT[3 + k] = 2.0f * s1 * dL_dc + s0 * dL_db;
      }
    }
    {
      float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        float w0 = __ldg(a.view + 4 * k), w1 = __ldg(a.view + 4 * k + 1), w2 = __ldg(a.view + 4 * k + 2);
        dJ00 += w0 * dT[k]; dJ02 += w2 * dT[k];
        dJ11 += w1 * dT[3 + k]; dJ12 += w2 * dT[3 + k];
      }
      float fx = (float)a.W / (2.0f * a.tanfovx), fy = (float)a.H / (2.0f * a.tanfovy);
      float tx = c2.t[0], ty = c2.t[1], tz = c2.t[2];
      float tzi = 1.0f / tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
      float dtx = xg * (-fx * tz2 * dJ02);
      float dty = yg * (-fy * tz2 * dJ12);
      float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2.0f * fx * tx) * tz3 * dJ02 +
                  (2.0f * fy * ty) * tz3 * dJ12;
      dtz += gcol.w;  // depth channel
#pragma unroll
      for (int k = 0; k < 3; k++)
        dmean[k] += __ldg(a.view + 4 * k) * dtx + __ldg(a.view + 4 * k + 1) * dty + __ldg(a.view + 4 * k + 2) * dtz;
    }
    // ---- mean2D -> mean3D ----
    {
      float ph[4];
      xform4x4(a.proj, p, ph);
      float mw = 1.0f / (ph[3] + 0.0000001f);
      float mul1 = ph[0] * mw * mw, mul2 = ph[1] * mw * mw;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        float p0 = __ldg(a.proj + 4 * k), p1 = __ldg(a.proj + 4 * k + 1), p3 = __ldg(a.proj + 4 * k + 3);
        dmean[k] += (p0 * mw - p3 * mul1) * gm.x + (p1 * mw - p3 * mul2) * gm.y;
      }
    }
    // ---- colour ----
    if (colors_precomp) {
      if (dL_dcolors_precomp) {
        dL_dcolors_precomp[3 * si] = gcol.x; dL_dcolors_precomp[3 * si + 1] = gcol.y; dL_dcolors_precomp[3 * si + 2] = gcol.z;
      }
    } else {
      float dv[3] = {p[0] - __ldg(a.campos), p[1] - __ldg(a.campos + 1), p[2] - __ldg(a.campos + 2)};
      float len = sqrtf(dv[0] * dv[0] + dv[1] * dv[1] + dv[2] * dv[2]);
      float inv = 1.0f / len;
      float d[3] = {dv[0] * inv, dv[1] * inv, dv[2] * inv};
      float b[16];
      sh_basis(a.sh_degree, d, b);
      uint8_t cb = clamped[i];
      float dRGB[3] = {(cb & 1) ? 0.f : gcol.x, (cb & 2) ? 0.f : gcol.y, (cb & 4) ? 0.f : gcol.z};
      const float* sh = shs + si * (kSplit ? 1 : a.M) * 3;
      float sdot[16];
      if (kSplit || (a.M == 16 && (((uintptr_t)shs) & 15) == 0)) {
        float v[48];
        if (kSplit) {
          const float* row = &s_rest[threadIdx.x >> 5][(threadIdx.x & 31) * kShRestRow];
          v[0] = sh[0]; v[1] = sh[1]; v[2] = sh[2];
#pragma unroll
          for (int e = 0; e < kShRestRow; e++) v[3 + e] = (e < 3 * (nb - 1)) ? row[e] : 0.f;
        } else {   // vector loads of the 192-byte coefficient block
          const float4* s4 = reinterpret_cast<const float4*>(sh);
#pragma unroll
          for (int q = 0; q < 12; q++) {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (4 * q < 3 * nb) t = s4[q];
            v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
          }
        }
#pragma unroll
        for (int k = 0; k < 16; k++) {
          sdot[k] = 0.f;
          if (k < nb) {
            sdot[k] = v[3 * k] * dRGB[0] + v[3 * k + 1] * dRGB[1] + v[3 * k + 2] * dRGB[2];
            shb[k] = b[k];
          }
        }
      } else {
#pragma unroll
        for (int k = 0; k < 16; k++) {
          sdot[k] = 0.f;
          if (k < nb && k < a.M) {
            sdot[k] = sh[3 * k] * dRGB[0] + sh[3 * k + 1] * dRGB[1] + sh[3 * k + 2] * dRGB[2];
            shb[k] = b[k];
          }
        }
      }
      shg[0] = dRGB[0]; shg[1] = dRGB[1]; shg[2] = dRGB[2];
      float x = d[0], y = d[1], z = d[2];
      float ddx = 0.f, ddy = 0.f, ddz = 0.f;
      if (a.sh_degree > 0) {
        ddy += -D3G_SH_C1 * sdot[1]; ddz += D3G_SH_C1 * sdot[2]; ddx += -D3G_SH_C1 * sdot[3];
        if (a.sh_degree > 1) {
          float xx = x * x, yy = y * y, zz = z * z;
          ddx += kSHC2[0] * y * sdot[4]; ddy += kSHC2[0] * x * sdot[4];
          ddy += kSHC2[1] * z * sdot[5]; ddz += kSHC2[1] * y * sdot[5];
          ddx += kSHC2[2] * -2.0f * x * sdot[6]; ddy += kSHC2[2] * -2.0f * y * sdot[6]; ddz += kSHC2[2] * 4.0f * z * sdot[6];
          ddx += kSHC2[3] * z * sdot[7]; ddz += kSHC2[3] * x * sdot[7];
          ddx += kSHC2[4] * 2.0f * x * sdot[8]; ddy += kSHC2[4] * -2.0f * y * sdot[8];
          if (a.sh_degree > 2) {
            ddx += kSHC3[0] * 6.0f * x * y * sdot[9]; ddy += kSHC3[0] * (3.0f * xx - 3.0f * yy) * sdot[9];
            ddx += kSHC3[1] * y * z * sdot[10]; ddy += kSHC3[1] * x * z * sdot[10]; ddz += kSHC3[1] * x * y * sdot[10];
            ddx += kSHC3[2] * -2.0f * x * y * sdot[11]; ddy += kSHC3[2] * (4.0f * zz - xx - 3.0f * yy) * sdot[11];
            ddz += kSHC3[2] * 8.0f * y * z * sdot[11];
            ddx += kSHC3[3] * -6.0f * x * z * sdot[12]; ddy += kSHC3[3] * -6.0f * y * z * sdot[12];
            ddz += kSHC3[3] * (6.0f * zz - 3.0f * xx - 3.0f * yy) * sdot[12];
            ddx += kSHC3[4] * (4.0f * zz - 3.0f * xx - yy) * sdot[13]; ddy += kSHC3[4] * -2.0f * x * y * sdot[13];
            ddz += kSHC3[4] * 8.0f * x * z * sdot[13];
            ddx += kSHC3[5] * 2.0f * x * z * sdot[14]; ddy += kSHC3[5] * -2.0f * y * z * sdot[14];
            ddz += kSHC3[5] * (xx - yy) * sdot[14];
            ddx += kSHC3[6] * (3.0f * xx - 3.0f * yy) * sdot[15]; ddy += kSHC3[6] * -6.0f * x * y * sdot[15];
          }
        }
      }
      float dot = d[0] * ddx + d[1] * ddy + d[2] * ddz;
      dmean[0] += (ddx - d[0] * dot) * inv;
      dmean[1] += (ddy - d[1] * dot) * inv;
      dmean[2] += (ddz - d[2] * dot) * inv;
    }
    // ---- cov3D -> scale / rotation ----
    if (scales && rots) {
      float4 qv = *reinterpret_cast<const float4*>(rots + 4 * si);
      float q[4] = {qv.x, qv.y, qv.z, qv.w};
      float R[9];
      quat_to_R(q, R);
      float mod = a.scale_modifier;
      float s[3] = {mod * scales[3 * si], mod * scales[3 * si + 1], mod * scales[3 * si + 2]};
      float dS[9] = {dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                     0.5f * dcov[2], 0.5f * dcov[4], dcov[5]};
      float L[9], dL[9], dR[9];
#pragma unroll
      for (int r = 0; r < 3; r++)
#pragma unroll
        for (int k = 0; k < 3; k++) L[3 * r + k] = R[3 * r + k] * s[k];
#pragma unroll
      for (int r = 0; r < 3; r++)
#pragma unroll
        for (int k = 0; k < 3; k++)
          dL[3 * r + k] = 2.0f * (dS[3 * r] * L[k] + dS[3 * r + 1] * L[3 + k] + dS[3 * r + 2] * L[6 + k]);
#pragma unroll
      for (int k = 0; k < 3; k++) {
        float g = 0.f;
#pragma unroll
        for (int r = 0; r < 3; r++) { g += dL[3 * r + k] * R[3 * r + k]; dR[3 * r + k] = dL[3 * r + k] * s[k]; }
        dsc[k] = g * mod;
      }
      float r = q[0], x = q[1], y = q[2], z = q[3];
      dq[0] = 2.0f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
      dq[1] = 2.0f * (y * dR[1] + z * dR[2] + y * dR[3] - 2.0f * x * dR[4] - r * dR[5] + z * dR[6] + r * dR[7] - 2.0f * x * dR[8]);
      dq[2] = 2.0f * (-2.0f * y * dR[0] + x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] + z * dR[7] - 2.0f * y * dR[8]);
      dq[3] = 2.0f * (-2.0f * z * dR[0] - r * dR[1] + x * dR[2] + r * dR[3] - 2.0f * z * dR[4] + y * dR[5] + x * dR[6] + y * dR[7]);
    }
  }
  dL_dmeans3D[3 * si] = dmean[0]; dL_dmeans3D[3 * si + 1] = dmean[1]; dL_dmeans3D[3 * si + 2] = dmean[2];
  if (dL_dscales) { dL_dscales[3 * si] = dsc[0]; dL_dscales[3 * si + 1] = dsc[1]; dL_dscales[3 * si + 2] = dsc[2]; }
  if (dL_drots) { dL_drots[4 * si] = dq[0]; dL_drots[4 * si + 1] = dq[1]; dL_drots[4 * si + 2] = dq[2]; dL_drots[4 * si + 3] = dq[3]; }
  if (dL_dcov3D) {
#pragma unroll
    for (int k = 0; k < 6; k++) dL_dcov3D[6 * si + k] = dcov[k];
  }
  if (kSplit) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int i0 = i - lane;
    const int cnt = min(32, a.N - i0);
    const unsigned mask = cnt >= 32 ? 0xffffffffu : ((1u << cnt) - 1u);
    // (every lane has consumed its coefficient row — it is only ever read by its own lane — so the block is free)
    if (!__any_sync(mask, vis)) return;   // a warp of culled Gaussians adds nothing
    if (vis) {
#pragma unroll
      for (int c = 0; c < 3; c++)   // (no contraction of the product into the sum: bit-identical to autograd's add)
        dL_dsh[3 * si + c] = __fadd_rn(dL_dsh[3 * si + c], __fmul_rn(shb[0], shg[c]));
    }
    float* row = &s_rest[warp][lane * kShRestRow];
#pragma unroll
    for (int e = 0; e < kShRestRow; e++) row[e] = shb[1 + e / 3] * shg[e % 3];
    __syncwarp(mask);
    const int nf = cnt * kShRestRow;
    float* out = dL_dsh_rest + (size_t)i0 * kShRestRow;
    float4* o4 = reinterpret_cast<float4*>(out);
    const float4* s4 = reinterpret_cast<const float4*>(s_rest[warp]);
    for (int q = lane; q < (nf >> 2); q += cnt) {     // only lanes 0..cnt-1 exist in a partial warp
      float4 g = o4[q];
      const float4 t = s4[q];
      g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
      o4[q] = g;
    }
    for (int e = (nf & ~3) + lane; e < nf; e += cnt) out[e] += s_rest[warp][e];
  } else if (dL_dsh) {
    __shared__ float s_sh[256 / 32][32][25];   // one 24-float slice of 32 Gaussians per warp (+1 pad: no bank conflicts)
    // the lanes of this warp that hold a Gaussian (the last warp of the grid may be partial); named explicitly —
    // __activemask() would only report the lanes that happen to have reconverged after the branches above
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cnt = min(32, a.N - (i - lane));
    const unsigned mask = cnt >= 32 ? 0xffffffffu : ((1u << cnt) - 1u);
    const int row = 3 * a.M;                   // floats per Gaussian
    float* out = dL_dsh + (si - lane) * row;   // first Gaussian of this warp
#pragma unroll
    for (int h = 0; h < 2; h++) {              // coefficients 0..7 and 8..15 (register indices stay compile-time)
      const int base = 24 * h;
      if (base >= row) break;
      const int len = min(24, row - base);
#pragma unroll
      for (int j = 0; j < 24; j++) s_sh[warp][lane][j] = shb[(base + j) / 3] * shg[(base + j) % 3];
      __syncwarp(mask);
      for (int idx = lane; idx < cnt * len; idx += cnt) {   // only lanes 0..cnt-1 exist in a partial warp
        const int g = idx / len, j = idx - g * len;
        out[(size_t)g * row + base + j] = s_sh[warp][g][j];
      }
      __syncwarp(mask);
    }
    for (int e = 48; e < row; e++) dL_dsh[si * row + e] = 0.f;   // coefficients past degree 3 (M > 16) get no gradient
  }
}

__global__ void mark_visible_kernel(int N, const float* __restrict__ means3D,
                                    const float* __restrict__ view, uint8_t* __restrict__ present) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float p[3] = {means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]};
  float pv[3];
  xform4x3(view, p, pv);
  present[i] = pv[2] > 0.2f ? 1 : 0;
}

static int bit_length(uint32_t n) { int b = 0; while (n >> b) b++; return b; }

static int check_settings(const d3g_raster_settings* s) {
  if (!s) return set_err(D3G_ERR_INVALID, "settings is NULL");
  if (s->image_height < 0 || s->image_width < 0) return set_err(D3G_ERR_INVALID, "negative image size");
  if (!s->bg || !s->viewmatrix || !s->projmatrix || !s->campos)
    return set_err(D3G_ERR_INVALID, "bg/viewmatrix/projmatrix/campos must be device pointers");
  if (s->sh_degree < 0 || s->sh_degree > 3) return set_err(D3G_ERR_UNSUPPORTED, "sh_degree must be 0..3");
  return 0;
}

static RasterArgs make_args(const d3g_raster_settings* s, int N, int M) {
  RasterArgs a;
  a.N = N; a.H = s->image_height; a.W = s->image_width;
  a.gx = (a.W + kTile - 1) / kTile; a.gy = (a.H + kTile - 1) / kTile;
  a.sh_degree = s->sh_degree; a.M = M;
  a.densify_mode = s->densify_mode ? s->densify_mode : 2;
  a.binning_mode = s->binning_mode;
  a.tanfovx = s->tanfovx; a.tanfovy = s->tanfovy; a.scale_modifier = s->scale_modifier;
  a.bg = s->bg; a.view = s->viewmatrix; a.proj = s->projmatrix; a.campos = s->campos;
  return a;
}

}  // namespace d3g

using namespace d3g;

extern "C" {

int d3g_raster_geom_bytes(int32_t N, size_t* bytes) {
  if (!bytes || N < 0) return set_err(D3G_ERR_INVALID, "bad arguments");
  *bytes = geom_view(nullptr, N).total;
  return 0;
}
int d3g_raster_img_bytes(int32_t H, int32_t W, size_t* bytes) {
  if (!bytes || H < 0 || W < 0) return set_err(D3G_ERR_INVALID, "bad arguments");
  *bytes = img_view(nullptr, H, W).total;
  return 0;
}
int d3g_raster_bin_bytes(uint64_t cap, size_t* bytes) {
  if (!bytes) return set_err(D3G_ERR_INVALID, "bad arguments");
  if (cap > 0x7fffffffull) return set_err(D3G_ERR_CAPACITY, "pair capacity above 2^31-1");
  *bytes = bin_view(nullptr, cap).total;
  return 0;
}

}  // extern "C"

// sh_rest != NULL: `shs` is _features_dc [N,1,3] and `sh_rest` _features_rest [N,15,3] (the *_split entry points)
static int raster_preprocess_impl(const d3g_raster_settings* s, int32_t N, const float* means3D,
                                  const float* shs, const float* sh_rest, int32_t M, const float* colors_precomp,
                                  const float* opacities, const float* scales, const float* rotations,
                                  const float* cov3D_precomp, void* geom, int32_t* radii,
                                  uint32_t* num_rendered_host, void* stream_) {
  if (int e = check_settings(s)) return e;
  cudaStream_t stream = (cudaStream_t)stream_;
  // (empty tensors have NULL data pointers, so the either/or checks only apply when N > 0; the Python
  // mirror performs them on the tensor objects themselves)
  if (N > 0 && (shs == nullptr) == (colors_precomp == nullptr))
    return set_err(D3G_ERR_INVALID, "Please provide excatly one of either SHs or precomputed colors!");
  if (N > 0 && (((scales == nullptr || rotations == nullptr) && cov3D_precomp == nullptr) ||
                ((scales != nullptr || rotations != nullptr) && cov3D_precomp != nullptr)))
    return set_err(D3G_ERR_INVALID,
                   "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
  if (N < 0 || !geom || (N > 0 && (!means3D || !opacities || !radii)))
    return set_err(D3G_ERR_INVALID, "NULL tensor");
  if (shs && M < (s->sh_degree + 1) * (s->sh_degree + 1))
    return set_err(D3G_ERR_INVALID, "shs has fewer coefficients than sh_degree needs");
  GeomView g = geom_view(geom, N);
  RasterArgs a = make_args(s, N, M);
  if (N > 0) {
    { ProfScope _p(PROF_PREPROCESS_FWD, stream);
    const bool tile_mode = a.binning_mode == 0;
    if (tile_mode) D3G_CUDA_TRY(cudaMemsetAsync(g.header, 0, 8, stream));
    if (sh_rest)
      preprocess_fwd_kernel<true><<<(N + 255) / 256, 256, 0, stream>>>(
          a, means3D, shs, sh_rest, colors_precomp, opacities, scales, rotations, cov3D_precomp, radii, g.means2D,
          g.conic_opacity, g.rgbd, g.cov3D, g.tiles_touched, g.clamped, tile_mode ? g.header : nullptr);
    else
      preprocess_fwd_kernel<false><<<(N + 255) / 256, 256, 0, stream>>>(
          a, means3D, shs, nullptr, colors_precomp, opacities, scales, rotations, cov3D_precomp, radii, g.means2D,
          g.conic_opacity, g.rgbd, g.cov3D, g.tiles_touched, g.clamped, tile_mode ? g.header : nullptr);
    D3G_LAUNCH_CHECK(s->debug, stream); }
    if (a.binning_mode != 0) {   // the reference's own scheme needs the per-Gaussian offsets
      size_t tb = g.scan_temp_bytes;
      ProfScope _p(PROF_SCAN, stream);
      D3G_CUDA_TRY(cub::DeviceScan::InclusiveSum(g.scan_temp, tb, g.tiles_touched, g.offsets, N, stream));
      finish_scan_kernel<<<1, 1, 0, stream>>>(N, g.offsets, g.header);
      D3G_LAUNCH_CHECK(s->debug, stream);
    }
  } else {
    finish_scan_kernel<<<1, 1, 0, stream>>>(N, g.offsets, g.header);
    D3G_LAUNCH_CHECK(s->debug, stream);
  }
  if (num_rendered_host)
    D3G_CUDA_TRY(cudaMemcpyAsync(num_rendered_host, g.header, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  return 0;
}

static int check_split(int32_t N, const float* dc, const float* rest, int32_t M) {
  if (M != 16) return set_err(D3G_ERR_UNSUPPORTED, "split SH features need sh_coeffs = 16 (max_sh_degree 3)");
  if (N > 0 && (!dc || !rest)) return set_err(D3G_ERR_INVALID, "features_dc / features_rest is NULL");
  if (((uintptr_t)rest) & 15) return set_err(D3G_ERR_INVALID, "features_rest must be 16-byte aligned");
  return 0;
}

extern "C" {

int d3g_raster_preprocess(const d3g_raster_settings* s, int32_t N, const float* means3D,
                          const float* shs, int32_t M, const float* colors_precomp,
                          const float* opacities, const float* scales, const float* rotations,
                          const float* cov3D_precomp, void* geom, int32_t* radii,
                          uint32_t* num_rendered_host, void* stream_) {
  return raster_preprocess_impl(s, N, means3D, shs, nullptr, M, colors_precomp, opacities, scales, rotations,
                                cov3D_precomp, geom, radii, num_rendered_host, stream_);
}

int d3g_raster_preprocess_split(const d3g_raster_settings* s, int32_t N, const float* means3D,
                                const float* features_dc, const float* features_rest, int32_t M,
                                const float* opacities, const float* scales, const float* rotations,
                                const float* cov3D_precomp, void* geom, int32_t* radii,
                                uint32_t* num_rendered_host, void* stream_) {
  if (int e = check_split(N, features_dc, features_rest, M)) return e;
  return raster_preprocess_impl(s, N, means3D, features_dc, N > 0 ? features_rest : nullptr, M, nullptr, opacities,
                                scales, rotations, cov3D_precomp, geom, radii, num_rendered_host, stream_);
}

int d3g_raster_render(const d3g_raster_settings* s, int32_t N, uint64_t cap, void* geom, void* bin,
                      void* img, const int32_t* radii, float* out_color, float* out_depth,
                      void* stream_) {
  if (int e = check_settings(s)) return e;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!geom || !img || (cap > 0 && !bin) || !out_color || !out_depth)
    return set_err(D3G_ERR_INVALID, "NULL workspace/output");
  if (cap > 0x7fffffffull) return set_err(D3G_ERR_CAPACITY, "pair capacity above 2^31-1");
  GeomView g = geom_view(geom, N);
  BinView b = bin_view(bin, cap);
  ImgView im = img_view(img, s->image_height, s->image_width);
  RasterArgs a = make_args(s, N, 0);
  size_t tiles = (size_t)a.gx * a.gy;
  if (tiles == 0) return 0;
  if (a.binning_mode == 0 && N > 0) {
    if (tiles > 0x7fffffffu) return set_err(D3G_ERR_CAPACITY, "too many tiles");
    D3G_CUDA_TRY(cudaMemsetAsync(im.tile_count, 0, tiles * kTileCtrStride * sizeof(uint32_t), stream));
    { ProfScope _p(PROF_TILE_COUNT, stream);
      tile_count_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, a.gx, a.gy, g.means2D, radii, im.tile_count);
      D3G_LAUNCH_CHECK(s->debug, stream); }
    { ProfScope _p(PROF_TILE_SCAN, stream);
      tile_scan_kernel<<<1, 1024, 0, stream>>>((int)tiles, cap, im.tile_count, im.tile_cursor, im.ranges, g.header);
      D3G_LAUNCH_CHECK(s->debug, stream); }
    if (cap > 0) {
      { ProfScope _p(PROF_TILE_SCATTER, stream);
        tile_scatter_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, a.gx, a.gy, g.means2D, g.rgbd, radii, im.tile_cursor,
                                                               b.keys_unsorted, g.header);
        D3G_LAUNCH_CHECK(s->debug, stream); }
      { ProfScope _p(PROF_TILE_SORT, stream);
        // (the keys_sorted buffer doubles as the merge scratch of long segments: every element is finally rewritten
        // in place by the thread that read it)
        tile_sort_kernel<<<(unsigned)tiles, 256, kTileSortCap * sizeof(uint64_t), stream>>>(
            im.ranges, b.keys_unsorted, b.keys_sorted, b.keys_sorted, b.point_list);
        D3G_LAUNCH_CHECK(s->debug, stream); }
    }
  } else {
    D3G_CUDA_TRY(cudaMemsetAsync(im.ranges, 0, tiles * sizeof(uint2), stream));
  }
  if (a.binning_mode != 0 && cap > 0 && N > 0) {
    { ProfScope _p(PROF_DUPLICATE, stream);
    duplicate_with_keys_kernel<<<(N + 255) / 256, 256, 0, stream>>>(
        N, a.gx, a.gy, cap, g.means2D, g.rgbd, g.offsets, radii, b.keys_unsorted, b.vals_unsorted, g.header);
    D3G_LAUNCH_CHECK(s->debug, stream); }
    { ProfScope _p(PROF_PAD_KEYS, stream);
    pad_keys_kernel<<<(unsigned)((cap + 255) / 256), 256, 0, stream>>>(cap, g.header, b.keys_unsorted, b.vals_unsorted);
    D3G_LAUNCH_CHECK(s->debug, stream); }
    size_t tb = b.sort_temp_bytes;
    int end_bit = 32 + bit_length((uint32_t)tiles);
    { ProfScope _p(PROF_SORT, stream);
      D3G_CUDA_TRY(cub::DeviceRadixSort::SortPairs(b.sort_temp, tb, b.keys_unsorted, b.keys_sorted,
                                                   b.vals_unsorted, b.point_list, (int)cap, 0, end_bit, stream)); }
    { ProfScope _p(PROF_RANGES, stream);
    identify_tile_ranges_kernel<<<(unsigned)((cap + 255) / 256), 256, 0, stream>>>(cap, g.header, b.keys_sorted, im.ranges);
    D3G_LAUNCH_CHECK(s->debug, stream); }
  }
  dim3 grid(a.gx, a.gy);
  { ProfScope _p(PROF_RENDER_FWD, stream);
    render_fwd_kernel<<<grid, kBlock, 0, stream>>>(a.H, a.W, a.gx, a.bg, im.ranges, b.point_list, g.means2D,
                                                 g.conic_opacity, g.rgbd, out_color, out_depth, im.final_T,
                                                 im.n_contrib);
  D3G_LAUNCH_CHECK(s->debug, stream); }
  return 0;
}

}  // extern "C"

// sh_rest != NULL: split SH features; dL_dshs / dL_dsh_rest are then the accumulating .grad tensors of the two parameters
static int raster_backward_impl(const d3g_raster_settings* s, int32_t N, uint64_t cap, const float* means3D,
                                const float* shs, const float* sh_rest, int32_t M, const float* colors_precomp,
                                const float* scales, const float* rotations, const float* cov3D_precomp,
                                const int32_t* radii, void* geom, void* bin, void* img, const float* dL_dcolor,
                                const float* dL_ddepth, float* dL_dmeans3D, float* dL_dmeans2D,
                                float* dL_dmeans2D_densify, float* dL_dshs, float* dL_dsh_rest,
                                float* dL_dcolors_precomp, float* dL_dopacities, float* dL_dscales,
                                float* dL_drotations, float* dL_dcov3D_precomp, void* stream_) {
  if (int e = check_settings(s)) return e;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!geom || !img || (cap > 0 && !bin) || !dL_dcolor) return set_err(D3G_ERR_INVALID, "NULL workspace/cotangent");
  if (N > 0 && !dL_dmeans3D) return set_err(D3G_ERR_INVALID, "dL_dmeans3D is required");
  if (shs && !dL_dshs) return set_err(D3G_ERR_INVALID, "dL_dshs is required when shs is given");
  (void)cov3D_precomp;
  GeomView g = geom_view(geom, N);
  BinView b = bin_view(bin, cap);
  ImgView im = img_view(img, s->image_height, s->image_width);
  RasterArgs a = make_args(s, N, M);
  if (N == 0) return 0;
  // the three accumulator arrays are contiguous: one memset
  D3G_CUDA_TRY(cudaMemsetAsync(g.acc_mean, 0, (size_t)((char*)g.scan_temp - (char*)g.acc_mean), stream));
  if ((size_t)a.gx * a.gy > 0) {
    dim3 grid(a.gx, a.gy);
    { ProfScope _p(PROF_RENDER_BWD, stream);
#define D3G_RBWD(DEPTH, MODE)                                                                                   \
    render_bwd_kernel<DEPTH, MODE><<<grid, kBlock, 0, stream>>>(a.H, a.W, a.gx, a.bg, im.ranges, b.point_list,       \
                                                                g.means2D, g.conic_opacity, g.rgbd, im.final_T,     \
                                                                im.n_contrib, dL_dcolor, dL_ddepth, g.acc_mean,     \
                                                                g.acc_conic, g.acc_color)
    const int dm = a.densify_mode == 1 ? 1 : (a.densify_mode == 2 ? 2 : 3);
    if (dL_ddepth) { if (dm == 1) D3G_RBWD(true, 1); else if (dm == 2) D3G_RBWD(true, 2); else D3G_RBWD(true, 3); }
    else { if (dm == 1) D3G_RBWD(false, 1); else if (dm == 2) D3G_RBWD(false, 2); else D3G_RBWD(false, 3); }
#undef D3G_RBWD
    D3G_LAUNCH_CHECK(s->debug, stream); }
  }
  { ProfScope _p(PROF_PREPROCESS_BWD, stream);
    if (sh_rest)
      preprocess_bwd_kernel<true><<<(N + 255) / 256, 256, 0, stream>>>(
        a, means3D, shs, sh_rest, colors_precomp, scales, rotations, radii, g.cov3D, g.clamped, g.acc_mean, g.acc_conic,
        g.acc_color, dL_dmeans3D, dL_dmeans2D, dL_dmeans2D_densify, dL_dshs, dL_dsh_rest, dL_dcolors_precomp,
        dL_dopacities, dL_dscales, dL_drotations, dL_dcov3D_precomp);
    else
      preprocess_bwd_kernel<false><<<(N + 255) / 256, 256, 0, stream>>>(
        a, means3D, shs, nullptr, colors_precomp, scales, rotations, radii, g.cov3D, g.clamped, g.acc_mean, g.acc_conic,
        g.acc_color, dL_dmeans3D, dL_dmeans2D, dL_dmeans2D_densify, dL_dshs, nullptr, dL_dcolors_precomp,
        dL_dopacities, dL_dscales, dL_drotations, dL_dcov3D_precomp);
  D3G_LAUNCH_CHECK(s->debug, stream); }
  return 0;
}

extern "C" {

int d3g_raster_backward(const d3g_raster_settings* s, int32_t N, uint64_t cap, const float* means3D,
                        const float* shs, int32_t M, const float* colors_precomp, const float* scales,
                        const float* rotations, const float* cov3D_precomp, const int32_t* radii,
                        void* geom, void* bin, void* img, const float* dL_dcolor, const float* dL_ddepth,
                        float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dmeans2D_densify,
                        float* dL_dshs, float* dL_dcolors_precomp, float* dL_dopacities,
                        float* dL_dscales, float* dL_drotations, float* dL_dcov3D_precomp,
                        void* stream_) {
  return raster_backward_impl(s, N, cap, means3D, shs, nullptr, M, colors_precomp, scales, rotations, cov3D_precomp,
                              radii, geom, bin, img, dL_dcolor, dL_ddepth, dL_dmeans3D, dL_dmeans2D,
                              dL_dmeans2D_densify, dL_dshs, nullptr, dL_dcolors_precomp, dL_dopacities, dL_dscales,
                              dL_drotations, dL_dcov3D_precomp, stream_);
}

int d3g_raster_backward_split(const d3g_raster_settings* s, int32_t N, uint64_t cap, const float* means3D,
                              const float* features_dc, const float* features_rest, int32_t M,
                              const float* scales, const float* rotations, const float* cov3D_precomp,
                              const int32_t* radii, void* geom, void* bin, void* img, const float* dL_dcolor,
                              const float* dL_ddepth, float* dL_dmeans3D, float* dL_dmeans2D,
                              float* dL_dmeans2D_densify, float* grad_features_dc, float* grad_features_rest,
                              float* dL_dopacities, float* dL_dscales, float* dL_drotations,
                              float* dL_dcov3D_precomp, void* stream_) {
  if (int e = check_split(N, features_dc, features_rest, M)) return e;
  if (N > 0 && (!grad_features_dc || !grad_features_rest))
    return set_err(D3G_ERR_INVALID, "grad_features_dc / grad_features_rest is NULL");
  if (((uintptr_t)grad_features_rest) & 15) return set_err(D3G_ERR_INVALID, "grad_features_rest must be 16-byte aligned");
  return raster_backward_impl(s, N, cap, means3D, features_dc, features_rest, M, nullptr, scales, rotations,
                              cov3D_precomp, radii, geom, bin, img, dL_dcolor, dL_ddepth, dL_dmeans3D, dL_dmeans2D,
                              dL_dmeans2D_densify, grad_features_dc, grad_features_rest, nullptr, dL_dopacities,
                              dL_dscales, dL_drotations, dL_dcov3D_precomp, stream_);
}

int d3g_mark_visible(int32_t N, const float* means3D, const float* viewmatrix, uint8_t* present,
                     void* stream_) {
  if (N < 0 || (N > 0 && (!means3D || !viewmatrix || !present))) return set_err(D3G_ERR_INVALID, "bad arguments");
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N > 0) {
    mark_visible_kernel<<<(N + 255) / 256, 256, 0, stream>>>(N, means3D, viewmatrix, present);
    D3G_LAUNCH_CHECK(0, stream);
  }
  return 0;
}

int d3g_raster_read_header(const void* geom, uint32_t header_host[2], void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!geom || !header_host) return set_err(D3G_ERR_INVALID, "bad arguments");
  D3G_CUDA_TRY(cudaMemcpyAsync(header_host, geom, 8, cudaMemcpyDeviceToHost, stream));
  D3G_CUDA_TRY(cudaStreamSynchronize(stream));
  return 0;
}

int d3g_raster_header_async(const void* geom, uint32_t* header_pinned, void* stream_) {
  if (!geom || !header_pinned) return set_err(D3G_ERR_INVALID, "bad arguments");
  D3G_CUDA_TRY(cudaMemcpyAsync(header_pinned, geom, 8, cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
  return 0;
}

int d3g_raster_read_state(int32_t which, int32_t N, int32_t H, int32_t W, uint64_t cap, const void* geom,
                          const void* bin, const void* img, void* dst, size_t dst_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GeomView g = geom_view((void*)geom, N);
  BinView b = bin_view((void*)bin, cap);
  ImgView im = img_view((void*)img, H, W);
  size_t tiles = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
  size_t P = (size_t)H * W, n = (size_t)N;
  const void* src = nullptr;
  size_t bytes = 0;
  switch (which) {
    case 0: src = g.means2D; bytes = n * 8; break;
    case 1: src = nullptr; bytes = n * 4; break;  // depths: strided .w of rgbd, handled below
    case 2: src = g.conic_opacity; bytes = n * 16; break;
    case 3: src = g.rgbd; bytes = n * 16; break;  // rgbd records (r,g,b,depth)
    case 4: src = g.cov3D; bytes = n * 24; break;
    case 5: src = g.tiles_touched; bytes = n * 4; break;
    case 6: src = g.offsets; bytes = n * 4; break;
    case 7: src = g.clamped; bytes = n; break;    // bitmask per Gaussian (bit c = channel c clamped)
    case 8: src = b.keys_sorted; bytes = (size_t)cap * 8; break;
    case 9: src = b.point_list; bytes = (size_t)cap * 4; break;
    case 10: src = im.ranges; bytes = tiles * 8; break;
    case 11: src = im.final_T; bytes = P * 4; break;
    case 12: src = im.n_contrib; bytes = P * 4; break;
    case 13: src = b.keys_unsorted; bytes = (size_t)cap * 8; break;
    case 14: src = b.vals_unsorted; bytes = (size_t)cap * 4; break;
    default: return set_err(D3G_ERR_INVALID, "unknown state id %d", which);
  }
  if (dst_bytes < bytes) return set_err(D3G_ERR_CAPACITY, "destination too small (%zu < %zu)", dst_bytes, bytes);
  if (bytes == 0) return 0;
  if (which == 1) {
    D3G_CUDA_TRY(cudaMemcpy2DAsync(dst, 4, (const char*)g.rgbd + 12, 16, 4, n, cudaMemcpyDeviceToHost, stream));
  } else {
    if ((which >= 8 && which <= 9) || which >= 13) { if (!bin) return set_err(D3G_ERR_INVALID, "bin is NULL"); }
    D3G_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream));
  }
  D3G_CUDA_TRY(cudaStreamSynchronize(stream));
  return 0;
}

}  // extern "C"
