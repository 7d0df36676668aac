/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Never imported, linked or executed by the product path.
 *
 * CPU restatement of the tile-based differentiable Gaussian rasterizer that the reference calls at
 * gaussian_renderer/__init__.py:53-68 (settings) and :115-124 (invocation).  The rasterizer source
 * itself (submodules/depth-diff-gaussian-rasterization, ingra14m/diff-gaussian-rasterization-extentions
 * @ d595eacabe86334988cd4273e0aaa07858b5ad2e, branch filter-norm) is NOT in /root/reference
 * (.SUBMODULES.json:9-15, status "missing"), so this file restates the published 3DGS rasterizer
 * algorithm as specified in SURVEY.md Appendix A, anchored on the reference call sites and on the
 * in-tree cross-checks:  SH polynomial = utils/sh_utils.py:57-112, cov3D = scene/gaussian_model.py:28-32
 * + utils/general_utils.py:130-163, matrix conventions = scene/cameras.py:55-61.
 *
 * PARITY UNPINNED: the reference ships no tests / golden vectors for this path and its
 * CUDA extension cannot be built or run here.  What IS pinned: analytic gradients vs. fp64 finite
 * differences of this file's own forward (tests/test_oracle_raster.py) and SH vs. the reference's
 * eval_sh (tests/golden/sh_eval.npz).
 *
 * This header is included twice by raster_ref.c: REAL=float (the bit-exact spec: every floating-point
 * operation is written out with explicit rounding points, FMA(a,b,c) = single-rounded a*b+c, so that
 * the CUDA kernels can reproduce radii / tiles / keys / ranges / image bit for bit) and REAL=double
 * (same formulas, used for finite-difference gradient checks).
 *
 * Compile with -ffp-contract=off so the compiler never fuses what is written as separate ops.
 */

#ifndef REAL
#error "include from raster_ref.c"
#endif

typedef struct {
  int32_t H, W;
  REAL tanfovx, tanfovy;
  REAL bg[3];
  REAL scale_modifier;
  REAL view[16]; /* world_view_transform, flat, as stored by scene/cameras.py:55 (transposed) */
  REAL proj[16]; /* full_proj_transform, flat (scene/cameras.py:59-60) */
  REAL campos[3];
  int32_t sh_degree; /* active degree 0..3 */
  int32_t M;         /* coefficients stored per Gaussian, (max_degree+1)^2 */
  int32_t densify_mode; /* semantics of dL/dmeans2D_densify, SURVEY.md §8(c): 1=H1 same as mean2D,
                           2=H2 component-wise sum of |per-pixel term|, 3=H3 sum of per-pixel 2-norm in .x */
} SFX(cfg);

/* ---- explicit-op helpers ------------------------------------------------------------------- */
static inline REAL SFX(rmin)(REAL a, REAL b) { return a < b ? a : b; }
static inline REAL SFX(rmax)(REAL a, REAL b) { return a > b ? a : b; }

/* Saturating float->int truncation (what cvt.rzi.s32.f32 does on the GPU; plain C cast is UB). */
static inline int32_t SFX(sat_int)(REAL f) {
  if (f != f) return 0;
  if (f >= (REAL)2147483648.0) return INT32_MAX;
  if (f <= (REAL)-2147483648.0) return INT32_MIN;
  return (int32_t)f;
}

/* Deterministic exp for x <= 0 (the only domain the blend uses).  IEEE-only operations, identical on
 * CPU and GPU: n = rint(x*log2e) by the magic-number trick, one-constant Cody-Waite reduction
 * (|n| <= ~130 here; for the |x|<6 range that passes the alpha test the reduction error is <1e-7),
 * degree-6 Taylor/Horner for e^r on |r|<=0.3466, exponent insertion by integer add.  Max relative
 * error vs. true exp on [-6,0]: ~2e-7.  Inputs below -87 return 0. */
static inline REAL SFX(det_exp)(REAL x) {
#if REAL_IS_FLOAT
  if (x < -87.0f) return 0.0f;
  const float magic = 12582912.0f; /* 1.5 * 2^23 */
  float t = FMA(x, 1.4426950408889634f, magic);
  float n = t - magic;
  float r = FMA(n, -0.6931471805599453f, x);
  float p = 1.0f / 720.0f;
  p = FMA(p, r, 1.0f / 120.0f);
  p = FMA(p, r, 1.0f / 24.0f);
  p = FMA(p, r, 1.0f / 6.0f);
  p = FMA(p, r, 0.5f);
  p = FMA(p, r, 1.0f);
  p = FMA(p, r, 1.0f);
  uint32_t tb, pb;
  memcpy(&tb, &t, 4);
  memcpy(&pb, &p, 4);
  pb += tb << 23;
  memcpy(&p, &pb, 4);
  return p;
#else
  return exp(x);
#endif
}

static inline void SFX(xform4x3)(const REAL* m, const REAL* p, REAL* o) {
  for (int a = 0; a < 3; a++) {
    REAL t = m[a] * p[0];
    t = FMA(m[4 + a], p[1], t);
    t = FMA(m[8 + a], p[2], t);
    o[a] = t + m[12 + a];
  }
}
static inline void SFX(xform4x4)(const REAL* m, const REAL* p, REAL* o) {
  for (int a = 0; a < 4; a++) {
    REAL t = m[a] * p[0];
    t = FMA(m[4 + a], p[1], t);
    t = FMA(m[8 + a], p[2], t);
    o[a] = t + m[12 + a];
  }
}

/* ndc2Pix evaluated in double, result rounded to REAL. */
static inline REAL SFX(ndc2pix)(REAL v, int S) {
  double t = (double)v + 1.0;
  double u = fma(t, (double)S, -1.0);
  return (REAL)(u * 0.5);
}

static inline void SFX(get_rect)(REAL px, REAL py, int radius, int gx, int gy, int* mnx, int* mny,
                                 int* mxx, int* mxy) {
  REAL r = (REAL)radius;
  int a = SFX(sat_int)((px - r) / (REAL)16.0);
  int b = SFX(sat_int)((py - r) / (REAL)16.0);
  int c = SFX(sat_int)((((px + r) + (REAL)16.0) - (REAL)1.0) / (REAL)16.0);
  int d = SFX(sat_int)((((py + r) + (REAL)16.0) - (REAL)1.0) / (REAL)16.0);
  *mnx = a < 0 ? 0 : (a > gx ? gx : a);
  *mny = b < 0 ? 0 : (b > gy ? gy : b);
  *mxx = c < 0 ? 0 : (c > gx ? gx : c);
  *mxy = d < 0 ? 0 : (d > gy ? gy : d);
}

/* Rotation (row-major standard rotation matrix of the UN-normalised quaternion (r,x,y,z)) times
 * diag(scale): L[a][k] = R[a][k]*s[k];  Sigma = L L^T.  Cross-check: scene/gaussian_model.py:28-32. */
static inline void SFX(quat_to_R)(const REAL* q, REAL R[9]) {
  REAL r = q[0], x = q[1], y = q[2], z = q[3];
  REAL xx = x * x, yy = y * y, zz = z * z;
  REAL xy = x * y, xz = x * z, yz = y * z, rx = r * x, ry = r * y, rz = r * z;
  R[0] = (REAL)1.0 - (REAL)2.0 * (yy + zz);
  R[1] = (REAL)2.0 * (xy - rz);
  R[2] = (REAL)2.0 * (xz + ry);
  R[3] = (REAL)2.0 * (xy + rz);
  R[4] = (REAL)1.0 - (REAL)2.0 * (xx + zz);
  R[5] = (REAL)2.0 * (yz - rx);
  R[6] = (REAL)2.0 * (xz - ry);
  R[7] = (REAL)2.0 * (yz + rx);
  R[8] = (REAL)1.0 - (REAL)2.0 * (xx + yy);
}

static inline void SFX(compute_cov3d)(const REAL* scale, REAL mod, const REAL* q, REAL* cov6) {
  REAL R[9], L[9];
  SFX(quat_to_R)(q, R);
  REAL s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
  for (int a = 0; a < 3; a++)
    for (int k = 0; k < 3; k++) L[3 * a + k] = R[3 * a + k] * s[k];
  int o = 0;
  for (int a = 0; a < 3; a++)
    for (int b = a; b < 3; b++) {
      REAL t = L[3 * a] * L[3 * b];
      t = FMA(L[3 * a + 1], L[3 * b + 1], t);
      t = FMA(L[3 * a + 2], L[3 * b + 2], t);
      cov6[o++] = t;
    }
}

/* EWA projection (App. A.1 step 5).  Outputs a,b,c (with +0.3 low-pass) and, for the backward pass,
 * the 2x3 matrix T = J*Wm and whether the frustum clamp was active. */
static inline void SFX(compute_cov2d)(const SFX(cfg) * c, const REAL* pv, const REAL* cov6, REAL* A,
                                      REAL* B, REAL* C, REAL T[6], int* clampx, int* clampy,
                                      REAL tclamped[3]) {
  REAL limx = (REAL)1.3 * c->tanfovx, limy = (REAL)1.3 * c->tanfovy;
  REAL tz = pv[2];
  REAL txtz = pv[0] / tz, tytz = pv[1] / tz;
  *clampx = (txtz < -limx) || (txtz > limx);
  *clampy = (tytz < -limy) || (tytz > limy);
  REAL tx = SFX(rmin)(limx, SFX(rmax)(-limx, txtz)) * tz;
  REAL ty = SFX(rmin)(limy, SFX(rmax)(-limy, tytz)) * tz;
  tclamped[0] = tx; tclamped[1] = ty; tclamped[2] = tz;
  REAL fx = (REAL)c->W / ((REAL)2.0 * c->tanfovx);
  REAL fy = (REAL)c->H / ((REAL)2.0 * c->tanfovy);
  REAL tz2 = tz * tz;
  REAL J00 = fx / tz, J02 = -(fx * tx) / tz2;
  REAL J11 = fy / tz, J12 = -(fy * ty) / tz2;
  const REAL* m = c->view;
  for (int k = 0; k < 3; k++) {
    /* Wm[a][k] = m[4*k + a] : view-space axis a as a function of world axis k */
    REAL w0 = m[4 * k + 0], w1 = m[4 * k + 1], w2 = m[4 * k + 2];
    T[k] = FMA(J02, w2, J00 * w0);
    T[3 + k] = FMA(J12, w2, J11 * w1);
  }
  REAL S[9] = {cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]};
  REAL v0[3], v1[3];
  for (int a = 0; a < 3; a++) {
    REAL t = S[3 * a] * T[0];
    t = FMA(S[3 * a + 1], T[1], t);
    v0[a] = FMA(S[3 * a + 2], T[2], t);
    REAL u = S[3 * a] * T[3];
    u = FMA(S[3 * a + 1], T[4], u);
    v1[a] = FMA(S[3 * a + 2], T[5], u);
  }
  REAL c00 = FMA(T[2], v0[2], FMA(T[1], v0[1], T[0] * v0[0]));
  REAL c01 = FMA(T[5], v0[2], FMA(T[4], v0[1], T[3] * v0[0]));
  REAL c11 = FMA(T[5], v1[2], FMA(T[4], v1[1], T[3] * v1[0]));
  *A = c00 + (REAL)0.3;
  *B = c01;
  *C = c11 + (REAL)0.3;
}

#define SH_C0 ((REAL)0.28209479177387814)
#define SH_C1 ((REAL)0.4886025119029199)
static const double SFX(SH_C2)[5] = {1.0925484305920792, -1.0925484305920792, 0.31539156525252005,
                                     -1.0925484305920792, 0.5462742152960396};
static const double SFX(SH_C3)[7] = {-0.5900435899266435, 2.890611442640554, -0.4570457994644658,
                                     0.3731763325901154, -0.4570457994644658, 1.445305721320277,
                                     -0.5900435899266435};

/* SH basis values b[0..15] for unit direction d (polynomial of utils/sh_utils.py:74-100; the colour is
 * sum_k b[k]*sh[k]).  Pinned op order: basis first, then a left-to-right FMA chain over k. */
static inline void SFX(sh_basis)(int deg, const REAL* d, REAL* b) {
  REAL x = d[0], y = d[1], z = d[2];
  b[0] = SH_C0;
  if (deg > 0) {
    b[1] = -(SH_C1 * y);
    b[2] = SH_C1 * z;
    b[3] = -(SH_C1 * x);
    if (deg > 1) {
      REAL xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = (REAL)SFX(SH_C2)[0] * xy;
      b[5] = (REAL)SFX(SH_C2)[1] * yz;
      b[6] = (REAL)SFX(SH_C2)[2] * ((((REAL)2.0 * zz) - xx) - yy);
      b[7] = (REAL)SFX(SH_C2)[3] * xz;
      b[8] = (REAL)SFX(SH_C2)[4] * (xx - yy);
      if (deg > 2) {
        b[9] = ((REAL)SFX(SH_C3)[0] * y) * (((REAL)3.0 * xx) - yy);
        b[10] = ((REAL)SFX(SH_C3)[1] * xy) * z;
        b[11] = ((REAL)SFX(SH_C3)[2] * y) * ((((REAL)4.0 * zz) - xx) - yy);
        b[12] = ((REAL)SFX(SH_C3)[3] * z) * ((((REAL)2.0 * zz) - ((REAL)3.0 * xx)) - ((REAL)3.0 * yy));
        b[13] = ((REAL)SFX(SH_C3)[4] * x) * ((((REAL)4.0 * zz) - xx) - yy);
        b[14] = ((REAL)SFX(SH_C3)[5] * z) * (xx - yy);
        b[15] = ((REAL)SFX(SH_C3)[6] * x) * (xx - ((REAL)3.0 * yy));
      }
    }
  }
}

/* ---- stage 1: preprocess (App. A.1) ---------------------------------------------------------- */
void SFX(preprocess_fwd)(const SFX(cfg) * c, int N, const REAL* means3D, const REAL* shs,
                         const REAL* colors_precomp, const REAL* opacities, const REAL* scales,
                         const REAL* rots, const REAL* cov3D_precomp, int32_t* radii, REAL* means2D,
                         REAL* depths, REAL* cov3D, REAL* rgb, REAL* conic_opacity,
                         uint32_t* tiles_touched, uint8_t* clamped) {
  int gx = (c->W + 15) / 16, gy = (c->H + 15) / 16;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < N; i++) {
    radii[i] = 0;
    tiles_touched[i] = 0;
    means2D[2 * i] = means2D[2 * i + 1] = 0;
    depths[i] = 0;
    for (int k = 0; k < 6; k++) cov3D[6 * i + k] = 0;
    for (int k = 0; k < 3; k++) { rgb[3 * i + k] = 0; clamped[3 * i + k] = 0; }
    for (int k = 0; k < 4; k++) conic_opacity[4 * i + k] = 0;
    const REAL* p = means3D + 3 * i;
    REAL pv[3], ph[4];
    SFX(xform4x3)(c->view, p, pv);
    if (!(pv[2] > (REAL)0.2)) continue;
    SFX(xform4x4)(c->proj, p, ph);
    REAL pw = (REAL)1.0 / (ph[3] + (REAL)0.0000001);
    REAL pp[2] = {ph[0] * pw, ph[1] * pw};
    REAL cov6[6];
    if (cov3D_precomp) {
      for (int k = 0; k < 6; k++) cov6[k] = cov3D_precomp[6 * i + k];
    } else {
      SFX(compute_cov3d)(scales + 3 * i, c->scale_modifier, rots + 4 * i, cov6);
    }
    REAL A, B, C, T[6], tcl[3];
    int cx, cy;
    SFX(compute_cov2d)(c, pv, cov6, &A, &B, &C, T, &cx, &cy, tcl);
    REAL det = FMA(A, C, -(B * B));
    if (det == (REAL)0.0) continue;
    REAL det_inv = (REAL)1.0 / det;
    REAL conic[3] = {C * det_inv, -B * det_inv, A * det_inv};
    REAL mid = (REAL)0.5 * (A + C);
    REAL disc = SFX(rmax)((REAL)0.1, FMA(mid, mid, -det));
    REAL sq = SQRT(disc);
    REAL l1 = mid + sq, l2 = mid - sq;
    REAL rad_f = CEIL((REAL)3.0 * SQRT(SFX(rmax)(l1, l2)));
    int my_radius = SFX(sat_int)(rad_f);
    REAL px = SFX(ndc2pix)(pp[0], c->W), py = SFX(ndc2pix)(pp[1], c->H);
    int mnx, mny, mxx, mxy;
    SFX(get_rect)(px, py, my_radius, gx, gy, &mnx, &mny, &mxx, &mxy);
    if ((mxx - mnx) * (mxy - mny) == 0) continue;
    if (colors_precomp) {
      for (int k = 0; k < 3; k++) rgb[3 * i + k] = colors_precomp[3 * i + k];
    } else {
      REAL d[3] = {p[0] - c->campos[0], p[1] - c->campos[1], p[2] - c->campos[2]};
      REAL len = SQRT(FMA(d[2], d[2], FMA(d[1], d[1], d[0] * d[0])));
      d[0] = d[0] / len; d[1] = d[1] / len; d[2] = d[2] / len;
      REAL b[16];
      SFX(sh_basis)(c->sh_degree, d, b);
      int nb = (c->sh_degree + 1) * (c->sh_degree + 1);
      const REAL* sh = shs + (size_t)i * c->M * 3;
      for (int ch = 0; ch < 3; ch++) {
        REAL r = b[0] * sh[ch];
        for (int k = 1; k < nb; k++) r = FMA(b[k], sh[3 * k + ch], r);
        r = r + (REAL)0.5;
        clamped[3 * i + ch] = (r < (REAL)0.0);
        rgb[3 * i + ch] = SFX(rmax)(r, (REAL)0.0);
      }
    }
    for (int k = 0; k < 6; k++) cov3D[6 * i + k] = cov6[k];
    depths[i] = pv[2];
    radii[i] = my_radius;
    means2D[2 * i] = px;
    means2D[2 * i + 1] = py;
    conic_opacity[4 * i] = conic[0];
    conic_opacity[4 * i + 1] = conic[1];
    conic_opacity[4 * i + 2] = conic[2];
    conic_opacity[4 * i + 3] = opacities[i];
    tiles_touched[i] = (uint32_t)((mxx - mnx) * (mxy - mny));
  }
}

/* ---- stage 2: binning (App. A.2) ------------------------------------------------------------- */
/* returns R; offsets[N] = inclusive scan of tiles_touched */
uint64_t SFX(scan_tiles)(int N, const uint32_t* tiles_touched, uint32_t* offsets) {
  uint64_t s = 0;
  for (int i = 0; i < N; i++) { s += tiles_touched[i]; offsets[i] = (uint32_t)s; }
  return s;
}

void SFX(duplicate_with_keys)(const SFX(cfg) * c, int N, const REAL* means2D, const REAL* depths,
                              const uint32_t* offsets, const int32_t* radii, uint64_t* keys,
                              uint32_t* vals) {
  int gx = (c->W + 15) / 16, gy = (c->H + 15) / 16;
  for (int i = 0; i < N; i++) {
    if (radii[i] <= 0) continue;
    uint32_t off = i ? offsets[i - 1] : 0;
    int mnx, mny, mxx, mxy;
    SFX(get_rect)(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, &mnx, &mny, &mxx, &mxy);
    float df = (float)depths[i];
    uint32_t db;
    memcpy(&db, &df, 4);
    for (int y = mny; y < mxy; y++)
      for (int x = mnx; x < mxx; x++) {
        uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
        key = (key << 32) | db;
        keys[off] = key;
        vals[off] = (uint32_t)i;
        off++;
      }
  }
}

/* smallest b with (n >> b) == 0 */
static inline int SFX(bit_length)(uint32_t n) { int b = 0; while (n >> b) b++; return b; }

void SFX(identify_tile_ranges)(uint64_t R, const uint64_t* keys, int num_tiles, uint32_t* ranges) {
  memset(ranges, 0, (size_t)num_tiles * 2 * sizeof(uint32_t));
  for (uint64_t j = 0; j < R; j++) {
    uint32_t tile = (uint32_t)(keys[j] >> 32);
    if (j == 0) ranges[2 * tile] = 0;
    else {
      uint32_t prev = (uint32_t)(keys[j - 1] >> 32);
      if (tile != prev) { ranges[2 * prev + 1] = (uint32_t)j; ranges[2 * tile] = (uint32_t)j; }
    }
    if (j == R - 1) ranges[2 * tile + 1] = (uint32_t)R;
  }
}

/* ---- stage 3: render forward (App. A.3) ------------------------------------------------------ */
void SFX(render_fwd)(const SFX(cfg) * c, const uint32_t* ranges, const uint32_t* point_list,
                     const REAL* means2D, const REAL* rgb, const REAL* depths,
                     const REAL* conic_opacity, REAL* out_color, REAL* out_depth, REAL* final_T,
                     uint32_t* n_contrib) {
  int gx = (c->W + 15) / 16, gy = (c->H + 15) / 16;
  int H = c->H, W = c->W;
#pragma omp parallel for schedule(dynamic, 4)
  for (int tile = 0; tile < gx * gy; tile++) {
    int tx = tile % gx, ty = tile / gx;
    uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
    for (int ly = 0; ly < 16; ly++)
      for (int lx = 0; lx < 16; lx++) {
        int pxi = tx * 16 + lx, pyi = ty * 16 + ly;
        if (pxi >= W || pyi >= H) continue;
        REAL pfx = (REAL)pxi, pfy = (REAL)pyi;
        REAL T = 1, C0 = 0, C1 = 0, C2 = 0, D = 0;
        uint32_t contributor = 0, last = 0;
        for (uint32_t j = lo; j < hi; j++) {
          contributor++;
          uint32_t id = point_list[j];
          REAL dx = means2D[2 * id] - pfx, dy = means2D[2 * id + 1] - pfy;
          const REAL* co = conic_opacity + 4 * id;
          REAL q = FMA(co[2] * dy, dy, (co[0] * dx) * dx);
          REAL power = FMA((REAL)-0.5, q, -((co[1] * dx) * dy));
          if (power > (REAL)0.0) continue;
          REAL alpha = SFX(rmin)((REAL)0.99, co[3] * SFX(det_exp)(power));
          if (alpha < (REAL)(1.0 / 255.0)) continue;
          REAL test_T = T * ((REAL)1.0 - alpha);
          if (test_T < (REAL)0.0001) break;
          REAL w = alpha * T;
          C0 = FMA(rgb[3 * id], w, C0);
          C1 = FMA(rgb[3 * id + 1], w, C1);
          C2 = FMA(rgb[3 * id + 2], w, C2);
          D = FMA(depths[id], w, D);
          T = test_T;
          last = contributor;
        }
        size_t pix = (size_t)pyi * W + pxi;
        final_T[pix] = T;
        n_contrib[pix] = last;
        out_color[pix] = FMA(T, c->bg[0], C0);
        out_color[(size_t)H * W + pix] = FMA(T, c->bg[1], C1);
        out_color[(size_t)2 * H * W + pix] = FMA(T, c->bg[2], C2);
        out_depth[pix] = D;
      }
  }
}

/* ---- stage 4: render backward (App. A.4) ----------------------------------------------------- */
/* Outputs are accumulated (+=); caller zero-initialises.  dL_dconic has 3 entries per Gaussian
 * (xx, xy, yy).  dL_dz is the gradient w.r.t. the view-space depth used by the depth channel. */
void SFX(render_bwd)(const SFX(cfg) * c, const uint32_t* ranges, const uint32_t* point_list,
                     const REAL* means2D, const REAL* conic_opacity, const REAL* rgb,
                     const REAL* depths, const REAL* final_T, const uint32_t* n_contrib,
                     const REAL* dL_dpix, const REAL* dL_ddepthpix, REAL* dL_dmean2D,
                     REAL* dL_dmean2D_densify, REAL* dL_dconic, REAL* dL_dopacity, REAL* dL_dcolor,
                     REAL* dL_dz) {
  int gx = (c->W + 15) / 16, gy = (c->H + 15) / 16;
  int H = c->H, W = c->W;
  REAL ddelx_dx = (REAL)0.5 * (REAL)W, ddely_dy = (REAL)0.5 * (REAL)H;
  for (int tile = 0; tile < gx * gy; tile++) {
    int tx = tile % gx, ty = tile / gx;
    uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
    for (int ly = 0; ly < 16; ly++)
      for (int lx = 0; lx < 16; lx++) {
        int pxi = tx * 16 + lx, pyi = ty * 16 + ly;
        if (pxi >= W || pyi >= H) continue;
        size_t pix = (size_t)pyi * W + pxi;
        REAL pfx = (REAL)pxi, pfy = (REAL)pyi;
        REAL T_final = final_T[pix];
        REAL T = T_final;
        uint32_t last_contributor = n_contrib[pix];
        REAL dLp[3] = {dL_dpix[pix], dL_dpix[(size_t)H * W + pix], dL_dpix[(size_t)2 * H * W + pix]};
        REAL dLd = dL_ddepthpix ? dL_ddepthpix[pix] : (REAL)0.0;
        REAL accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0};
        REAL accum_drec = 0, last_depth = 0, last_alpha = 0;
        REAL bg_dot = c->bg[0] * dLp[0] + c->bg[1] * dLp[1] + c->bg[2] * dLp[2];
        if (hi <= lo) continue;
        for (uint32_t k = last_contributor; k-- > 0;) {
          uint32_t j = lo + k; /* k = contributor index - 1 */
          uint32_t id = point_list[j];
          REAL dx = means2D[2 * id] - pfx, dy = means2D[2 * id + 1] - pfy;
          const REAL* co = conic_opacity + 4 * id;
          REAL q = FMA(co[2] * dy, dy, (co[0] * dx) * dx);
          REAL power = FMA((REAL)-0.5, q, -((co[1] * dx) * dy));
          if (power > (REAL)0.0) continue;
          REAL G = SFX(det_exp)(power);
          REAL alpha = SFX(rmin)((REAL)0.99, co[3] * G);
          if (alpha < (REAL)(1.0 / 255.0)) continue;
          T = T / ((REAL)1.0 - alpha);
          REAL dch = alpha * T;
          REAL dL_dalpha = 0;
          for (int ch = 0; ch < 3; ch++) {
            REAL col = rgb[3 * id + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + ((REAL)1.0 - last_alpha) * accum_rec[ch];
            last_color[ch] = col;
            dL_dalpha += (col - accum_rec[ch]) * dLp[ch];
            dL_dcolor[3 * id + ch] += dch * dLp[ch];
          }
          {
            REAL dep = depths[id];
            accum_drec = last_alpha * last_depth + ((REAL)1.0 - last_alpha) * accum_drec;
            last_depth = dep;
            dL_dalpha += (dep - accum_drec) * dLd;
            dL_dz[id] += dch * dLd;
          }
          dL_dalpha *= T;
          last_alpha = alpha;
          dL_dalpha += (-T_final / ((REAL)1.0 - alpha)) * bg_dot;
          REAL dL_dG = co[3] * dL_dalpha;
          REAL gdx = G * dx, gdy = G * dy;
          REAL dG_ddelx = -gdx * co[0] - gdy * co[1];
          REAL dG_ddely = -gdy * co[2] - gdx * co[1];
          REAL mx = dL_dG * dG_ddelx * ddelx_dx, my = dL_dG * dG_ddely * ddely_dy;
          dL_dmean2D[2 * id] += mx;
          dL_dmean2D[2 * id + 1] += my;
          if (c->densify_mode == 1) {
            dL_dmean2D_densify[2 * id] += mx;
            dL_dmean2D_densify[2 * id + 1] += my;
          } else if (c->densify_mode == 2) {
            dL_dmean2D_densify[2 * id] += (mx < 0 ? -mx : mx);
            dL_dmean2D_densify[2 * id + 1] += (my < 0 ? -my : my);
          } else {
            dL_dmean2D_densify[2 * id] += SQRT(mx * mx + my * my);
          }
          dL_dconic[3 * id] += (REAL)-0.5 * gdx * dx * dL_dG;
          dL_dconic[3 * id + 1] += (REAL)-0.5 * gdx * dy * dL_dG;
          dL_dconic[3 * id + 2] += (REAL)-0.5 * gdy * dy * dL_dG;
          dL_dopacity[id] += G * dL_dalpha;
        }
      }
  }
}

/* ---- stage 5: preprocess backward (App. A.4) ------------------------------------------------- */
void SFX(preprocess_bwd)(const SFX(cfg) * c, int N, const REAL* means3D, const REAL* shs,
                         const REAL* colors_precomp, const REAL* scales, const REAL* rots,
                         const REAL* cov3D_precomp, const int32_t* radii, const REAL* cov3D,
                         const uint8_t* clamped, const REAL* dL_dmean2D, const REAL* dL_dconic,
                         const REAL* dL_dcolor, const REAL* dL_dz, REAL* dL_dmeans3D, REAL* dL_dsh,
                         REAL* dL_dcolors_precomp, REAL* dL_dscales, REAL* dL_drots,
                         REAL* dL_dcov3D) {
  (void)cov3D_precomp;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < N; i++) {
    if (!(radii[i] > 0)) continue;
    const REAL* p = means3D + 3 * i;
    REAL dmean[3] = {0, 0, 0};
    /* ---- cov2D backward ---- */
    REAL pv[3];
    SFX(xform4x3)(c->view, p, pv);
    const REAL* cov6 = cov3D + 6 * i;
    REAL A, B, C, T[6], tcl[3];
    int cx, cy;
    SFX(compute_cov2d)(c, pv, cov6, &A, &B, &C, T, &cx, &cy, tcl);
    REAL xg = cx ? (REAL)0.0 : (REAL)1.0, yg = cy ? (REAL)0.0 : (REAL)1.0;
    REAL denom = A * C - B * B;
    REAL d2 = (REAL)1.0 / (denom * denom + (REAL)0.0000001);
    REAL dcx = dL_dconic[3 * i], dcy = dL_dconic[3 * i + 1], dcz = dL_dconic[3 * i + 2];
    REAL dL_da = 0, dL_db = 0, dL_dc = 0;
    REAL dcov[6] = {0, 0, 0, 0, 0, 0};
    REAL dT[6] = {0, 0, 0, 0, 0, 0};
    if (d2 != (REAL)0.0) {
      dL_da = d2 * (-C * C * dcx + (REAL)2.0 * B * C * dcy + (denom - A * C) * dcz);
      dL_dc = d2 * (-A * A * dcz + (REAL)2.0 * A * B * dcy + (denom - A * C) * dcx);
      dL_db = d2 * (REAL)2.0 * (B * C * dcx - (denom + (REAL)2.0 * B * B) * dcy + A * B * dcz);
      /* cov2D = T Sigma T^T (2x2); a=c00, b=c01, c=c11.  d/dSigma_ab (upper-triangle storage,
       * off-diagonals represent both symmetric entries). */
      int idx[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
      for (int a = 0; a < 3; a++)
        for (int b = a; b < 3; b++) {
          REAL g = T[a] * T[b] * dL_da + T[3 + a] * T[3 + b] * dL_dc;
          if (a == b) g += T[a] * T[3 + a] * dL_db;
          else {
            g = (REAL)2.0 * g;
            g += (T[a] * T[3 + b] + T[b] * T[3 + a]) * dL_db;
          }
          dcov[idx[a][b]] = g;
        }
      /* gradient w.r.t. T: cov = T S T^T */
      REAL S[9] = {cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]};
      for (int k = 0; k < 3; k++) {
        REAL s0 = S[3 * k] * T[0] + S[3 * k + 1] * T[1] + S[3 * k + 2] * T[2];       /* (S T0)_k */
        REAL s1 = S[3 * k] * T[3] + S[3 * k + 1] * T[4] + S[3 * k + 2] * T[5];       /* (S T1)_k */
        dT[k] = (REAL)2.0 * s0 * dL_da + s1 * dL_db;
        dT[3 + k] = (REAL)2.0 * s1 * dL_dc + s0 * dL_db;
      }
    }
    {
      /* T[0][k] = J00*W0k + J02*W2k ; T[1][k] = J11*W1k + J12*W2k */
      const REAL* m = c->view;
      REAL dJ00 = 0, dJ02 = 0, dJ11 = 0, dJ12 = 0;
      for (int k = 0; k < 3; k++) {
        REAL w0 = m[4 * k + 0], w1 = m[4 * k + 1], w2 = m[4 * k + 2];
        dJ00 += w0 * dT[k];
        dJ02 += w2 * dT[k];
        dJ11 += w1 * dT[3 + k];
        dJ12 += w2 * dT[3 + k];
      }
      REAL fx = (REAL)c->W / ((REAL)2.0 * c->tanfovx);
      REAL fy = (REAL)c->H / ((REAL)2.0 * c->tanfovy);
      REAL tx = tcl[0], ty = tcl[1], tz = tcl[2];
      REAL tzi = (REAL)1.0 / tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
      REAL dtx = xg * (-fx * tz2 * dJ02);
      REAL dty = yg * (-fy * tz2 * dJ12);
      REAL dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + ((REAL)2.0 * fx * tx) * tz3 * dJ02 +
                 ((REAL)2.0 * fy * ty) * tz3 * dJ12;
      /* depth channel: d(depth)/d(mean) = third row of the view rotation */
      dtz += dL_dz[i];
      for (int k = 0; k < 3; k++)
        dmean[k] += m[4 * k + 0] * dtx + m[4 * k + 1] * dty + m[4 * k + 2] * dtz;
    }
    /* ---- mean2D -> mean3D ---- */
    {
      const REAL* pr = c->proj;
      REAL ph[4];
      SFX(xform4x4)(pr, p, ph);
      REAL mw = (REAL)1.0 / (ph[3] + (REAL)0.0000001);
      REAL mul1 = ph[0] * mw * mw, mul2 = ph[1] * mw * mw;
      REAL gx_ = dL_dmean2D[2 * i], gy_ = dL_dmean2D[2 * i + 1];
      for (int k = 0; k < 3; k++)
        dmean[k] += (pr[4 * k] * mw - pr[4 * k + 3] * mul1) * gx_ +
                    (pr[4 * k + 1] * mw - pr[4 * k + 3] * mul2) * gy_;
    }
    /* ---- colour ---- */
    if (colors_precomp) {
      for (int ch = 0; ch < 3; ch++) dL_dcolors_precomp[3 * i + ch] = dL_dcolor[3 * i + ch];
    } else {
      REAL dvec[3] = {p[0] - c->campos[0], p[1] - c->campos[1], p[2] - c->campos[2]};
      REAL len = SQRT(dvec[0] * dvec[0] + dvec[1] * dvec[1] + dvec[2] * dvec[2]);
      REAL d[3] = {dvec[0] / len, dvec[1] / len, dvec[2] / len};
      REAL b[16];
      SFX(sh_basis)(c->sh_degree, d, b);
      int nb = (c->sh_degree + 1) * (c->sh_degree + 1);
      const REAL* sh = shs + (size_t)i * c->M * 3;
      REAL* dsh = dL_dsh + (size_t)i * c->M * 3;
      REAL dRGB[3];
      for (int ch = 0; ch < 3; ch++) dRGB[ch] = clamped[3 * i + ch] ? (REAL)0.0 : dL_dcolor[3 * i + ch];
      for (int k = 0; k < nb; k++)
        for (int ch = 0; ch < 3; ch++) dsh[3 * k + ch] = b[k] * dRGB[ch];
      /* d(colour)/d(direction): analytic derivative of the basis */
      REAL x = d[0], y = d[1], z = d[2];
      REAL db[16][3];
      memset(db, 0, sizeof(db));
      if (c->sh_degree > 0) {
        db[1][1] = -SH_C1; db[2][2] = SH_C1; db[3][0] = -SH_C1;
        if (c->sh_degree > 1) {
          REAL xx = x * x, yy = y * y, zz = z * z;
          const double* C2 = SFX(SH_C2); const double* C3 = SFX(SH_C3);
          db[4][0] = (REAL)C2[0] * y; db[4][1] = (REAL)C2[0] * x;
          db[5][1] = (REAL)C2[1] * z; db[5][2] = (REAL)C2[1] * y;
          db[6][0] = (REAL)C2[2] * (REAL)-2.0 * x; db[6][1] = (REAL)C2[2] * (REAL)-2.0 * y;
          db[6][2] = (REAL)C2[2] * (REAL)4.0 * z;
          db[7][0] = (REAL)C2[3] * z; db[7][2] = (REAL)C2[3] * x;
          db[8][0] = (REAL)C2[4] * (REAL)2.0 * x; db[8][1] = (REAL)C2[4] * (REAL)-2.0 * y;
          if (c->sh_degree > 2) {
            db[9][0] = (REAL)C3[0] * (REAL)6.0 * x * y;
            db[9][1] = (REAL)C3[0] * ((REAL)3.0 * xx - (REAL)3.0 * yy);
            db[10][0] = (REAL)C3[1] * y * z; db[10][1] = (REAL)C3[1] * x * z; db[10][2] = (REAL)C3[1] * x * y;
            db[11][0] = (REAL)C3[2] * (REAL)-2.0 * x * y;
            db[11][1] = (REAL)C3[2] * ((REAL)4.0 * zz - xx - (REAL)3.0 * yy);
            db[11][2] = (REAL)C3[2] * (REAL)8.0 * y * z;
            db[12][0] = (REAL)C3[3] * (REAL)-6.0 * x * z; db[12][1] = (REAL)C3[3] * (REAL)-6.0 * y * z;
            db[12][2] = (REAL)C3[3] * ((REAL)6.0 * zz - (REAL)3.0 * xx - (REAL)3.0 * yy);
            db[13][0] = (REAL)C3[4] * ((REAL)4.0 * zz - (REAL)3.0 * xx - yy);
            db[13][1] = (REAL)C3[4] * (REAL)-2.0 * x * y;
            db[13][2] = (REAL)C3[4] * (REAL)8.0 * x * z;
            db[14][0] = (REAL)C3[5] * (REAL)2.0 * x * z; db[14][1] = (REAL)C3[5] * (REAL)-2.0 * y * z;
            db[14][2] = (REAL)C3[5] * (xx - yy);
            db[15][0] = (REAL)C3[6] * ((REAL)3.0 * xx - (REAL)3.0 * yy);
            db[15][1] = (REAL)C3[6] * (REAL)-6.0 * x * y;
          }
        }
      }
      REAL ddir[3] = {0, 0, 0};
      for (int k = 1; k < nb; k++)
        for (int ax = 0; ax < 3; ax++) {
          REAL s = sh[3 * k] * dRGB[0] + sh[3 * k + 1] * dRGB[1] + sh[3 * k + 2] * dRGB[2];
          ddir[ax] += db[k][ax] * s;
        }
      /* through the normalisation d = v/|v| */
      REAL inv = (REAL)1.0 / len;
      REAL dot = d[0] * ddir[0] + d[1] * ddir[1] + d[2] * ddir[2];
      for (int k = 0; k < 3; k++) dmean[k] += (ddir[k] - d[k] * dot) * inv;
    }
    for (int k = 0; k < 3; k++) dL_dmeans3D[3 * i + k] = dmean[k];
    /* ---- cov3D -> scale / rotation ---- */
    for (int k = 0; k < 6; k++) dL_dcov3D[6 * i + k] = dcov[k];
    if (scales && rots) {
      const REAL* q = rots + 4 * i;
      REAL R[9];
      SFX(quat_to_R)(q, R);
      REAL mod = c->scale_modifier;
      REAL s[3] = {mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2]};
      /* symmetric dSigma with off-diagonals halved (storage convention: off-diagonal entry = both) */
      REAL dS[9] = {dcov[0], (REAL)0.5 * dcov[1], (REAL)0.5 * dcov[2], (REAL)0.5 * dcov[1], dcov[3],
                    (REAL)0.5 * dcov[4], (REAL)0.5 * dcov[2], (REAL)0.5 * dcov[4], dcov[5]};
      /* Sigma = L L^T, L = R diag(s):  dL = 2 dS L */
      REAL L[9], dL[9];
      for (int a = 0; a < 3; a++) for (int k = 0; k < 3; k++) L[3 * a + k] = R[3 * a + k] * s[k];
      for (int a = 0; a < 3; a++)
        for (int k = 0; k < 3; k++)
          dL[3 * a + k] = (REAL)2.0 * (dS[3 * a] * L[k] + dS[3 * a + 1] * L[3 + k] + dS[3 * a + 2] * L[6 + k]);
      REAL dR[9];
      for (int k = 0; k < 3; k++) {
        REAL g = 0;
        for (int a = 0; a < 3; a++) { g += dL[3 * a + k] * R[3 * a + k]; dR[3 * a + k] = dL[3 * a + k] * s[k]; }
        dL_dscales[3 * i + k] = g * mod;
      }
      REAL r = q[0], x = q[1], y = q[2], z = q[3];
      REAL dq[4];
      dq[0] = (REAL)2.0 * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
      dq[1] = (REAL)2.0 * (y * dR[1] + z * dR[2] + y * dR[3] - (REAL)2.0 * x * dR[4] - r * dR[5] + z * dR[6] +
                           r * dR[7] - (REAL)2.0 * x * dR[8]);
      dq[2] = (REAL)2.0 * (-(REAL)2.0 * y * dR[0] + x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] +
                           z * dR[7] - (REAL)2.0 * y * dR[8]);
      dq[3] = (REAL)2.0 * (-(REAL)2.0 * z * dR[0] - r * dR[1] + x * dR[2] + r * dR[3] - (REAL)2.0 * z * dR[4] +
                           y * dR[5] + x * dR[6] + y * dR[7]);
      for (int k = 0; k < 4; k++) dL_drots[4 * i + k] = dq[k];
    }
  }
}
