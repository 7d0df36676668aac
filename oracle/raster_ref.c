/*
 * ORACLE — TEST INFRASTRUCTURE ONLY (see raster_ref_impl.h for provenance and the "parity unpinned"
 * statement).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load the library built from this file.
 *
 * Build (see oracle/Makefile):  gcc -O2 -fPIC -shared -ffp-contract=off -fopenmp raster_ref.c -lm
 *
 * Exports two instantiations of every stage: *_f32 (bit-exact spec) and *_f64 (gradient checks),
 * plus the stable 64-bit key sort and a whole-frame convenience entry used for CPU timing.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---------------- float instantiation ---------------- */
#define REAL float
#define REAL_IS_FLOAT 1
#define SFX(n) n##_f32
#define FMA(a, b, c) fmaf((a), (b), (c))
#define SQRT(a) sqrtf(a)
#define CEIL(a) ceilf(a)
#include "raster_ref_impl.h"
#undef REAL
#undef REAL_IS_FLOAT
#undef SFX
#undef FMA
#undef SQRT
#undef CEIL
#undef SH_C0
#undef SH_C1

/* ---------------- double instantiation ---------------- */
#define REAL double
#define REAL_IS_FLOAT 0
#define SFX(n) n##_f64
#define FMA(a, b, c) fma((a), (b), (c))
#define SQRT(a) sqrt(a)
#define CEIL(a) ceil(a)
#include "raster_ref_impl.h"
#undef REAL
#undef REAL_IS_FLOAT
#undef SFX
#undef FMA
#undef SQRT
#undef CEIL

/* Stable LSD radix sort of (key,value) pairs on key bits [0,end_bit) — the contract of the 64-bit
 * tile|depth sort: ties keep emission order. */
void sort_pairs_u64(uint64_t R, int end_bit, uint64_t* keys, uint32_t* vals) {
  if (R == 0) return;
  uint64_t* k2 = (uint64_t*)malloc(R * sizeof(uint64_t));
  uint32_t* v2 = (uint32_t*)malloc(R * sizeof(uint32_t));
  size_t* hist = (size_t*)malloc(65536 * sizeof(size_t));
  uint64_t *ka = keys, *kb = k2;
  uint32_t *va = vals, *vb = v2;
  for (int shift = 0; shift < end_bit; shift += 16) {
    int bits = end_bit - shift < 16 ? end_bit - shift : 16;
    uint64_t mask = (1ull << bits) - 1;
    memset(hist, 0, 65536 * sizeof(size_t));
    for (uint64_t i = 0; i < R; i++) hist[(ka[i] >> shift) & mask]++;
    size_t s = 0;
    for (int b = 0; b < 65536; b++) { size_t t = hist[b]; hist[b] = s; s += t; }
    for (uint64_t i = 0; i < R; i++) {
      size_t d = hist[(ka[i] >> shift) & mask]++;
      kb[d] = ka[i];
      vb[d] = va[i];
    }
    uint64_t* tk = ka; ka = kb; kb = tk;
    uint32_t* tv = va; va = vb; vb = tv;
  }
  if (ka != keys) {
    memcpy(keys, ka, R * sizeof(uint64_t));
    memcpy(vals, va, R * sizeof(uint32_t));
  }
  free(k2); free(v2); free(hist);
}

int sort_end_bit(int num_tiles) { return 32 + bit_length_f32((uint32_t)num_tiles); }

/* Whole-frame fp32 forward+backward on the CPU (used only to time the restated CPU baseline).
 * Returns R.  Caller provides all buffers; bin buffers must hold R_cap pairs. */
uint64_t frame_fwd_bwd_f32(const cfg_f32* c, int N, const float* means3D, const float* shs,
                           const float* opacities, const float* scales, const float* rots,
                           const float* dL_dpix, float* out_color, float* out_depth,
                           float* dL_dmeans3D, float* dL_dsh, float* dL_dopacity, float* dL_dscales,
                           float* dL_drots, uint64_t R_cap, int do_backward) {
  int gx = (c->W + 15) / 16, gy = (c->H + 15) / 16;
  size_t P = (size_t)c->H * c->W;
  int32_t* radii = (int32_t*)malloc(N * sizeof(int32_t));
  float* means2D = (float*)malloc((size_t)N * 2 * 4);
  float* depths = (float*)malloc((size_t)N * 4);
  float* cov3D = (float*)malloc((size_t)N * 6 * 4);
  float* rgb = (float*)malloc((size_t)N * 3 * 4);
  float* co = (float*)malloc((size_t)N * 4 * 4);
  uint32_t* tt = (uint32_t*)malloc((size_t)N * 4);
  uint32_t* offs = (uint32_t*)malloc((size_t)N * 4);
  uint8_t* clamped = (uint8_t*)malloc((size_t)N * 3);
  preprocess_fwd_f32(c, N, means3D, shs, NULL, opacities, scales, rots, NULL, radii, means2D, depths,
                     cov3D, rgb, co, tt, clamped);
  uint64_t R = scan_tiles_f32(N, tt, offs);
  if (R > R_cap) R_cap = R;
  uint64_t* keys = (uint64_t*)malloc((R + 1) * 8);
  uint32_t* vals = (uint32_t*)malloc((R + 1) * 4);
  uint32_t* ranges = (uint32_t*)malloc((size_t)gx * gy * 8);
  duplicate_with_keys_f32(c, N, means2D, depths, offs, radii, keys, vals);
  sort_pairs_u64(R, sort_end_bit(gx * gy), keys, vals);
  identify_tile_ranges_f32(R, keys, gx * gy, ranges);
  float* final_T = (float*)malloc(P * 4);
  uint32_t* ncontrib = (uint32_t*)malloc(P * 4);
  render_fwd_f32(c, ranges, vals, means2D, rgb, depths, co, out_color, out_depth, final_T, ncontrib);
  if (do_backward) {
    float* g2d = (float*)calloc((size_t)N * 2, 4);
    float* g2dd = (float*)calloc((size_t)N * 2, 4);
    float* gcon = (float*)calloc((size_t)N * 3, 4);
    float* gcol = (float*)calloc((size_t)N * 3, 4);
    float* gz = (float*)calloc((size_t)N, 4);
    float* gcov = (float*)calloc((size_t)N * 6, 4);
    memset(dL_dopacity, 0, (size_t)N * 4);
    memset(dL_dmeans3D, 0, (size_t)N * 12);
    memset(dL_dsh, 0, (size_t)N * c->M * 12);
    memset(dL_dscales, 0, (size_t)N * 12);
    memset(dL_drots, 0, (size_t)N * 16);
    render_bwd_f32(c, ranges, vals, means2D, co, rgb, depths, final_T, ncontrib, dL_dpix, NULL, g2d,
                   g2dd, gcon, dL_dopacity, gcol, gz);
    preprocess_bwd_f32(c, N, means3D, shs, NULL, scales, rots, NULL, radii, cov3D, clamped, g2d, gcon,
                       gcol, gz, dL_dmeans3D, dL_dsh, NULL, dL_dscales, dL_drots, gcov);
    free(g2d); free(g2dd); free(gcon); free(gcol); free(gz); free(gcov);
  }
  free(radii); free(means2D); free(depths); free(cov3D); free(rgb); free(co); free(tt); free(offs);
  free(clamped); free(keys); free(vals); free(ranges); free(final_T); free(ncontrib);
  return R;
}

/* direct access to the pinned exp (accuracy test) */
void det_exp_array_f32(int n, const float* x, float* y) {
  for (int i = 0; i < n; i++) y[i] = det_exp_f32(x[i]);
}
