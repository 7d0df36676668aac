/* ORACLE / TEST INFRASTRUCTURE ONLY — never imported by the product package.
 * CPU restatement of `distCUDA2` (submodules/simple-knn/spatial.cu:15-26 -> SimpleKNN::knn, simple_knn.cu:185-221):
 * for every point the mean of the three smallest squared distances to the points at the other indices.  The
 * reference's Morton sort + box pruning (simple_knn.cu:139-183) is an exact search, so a brute-force scan gives the
 * same multiset; `updateKBest<3>` (simple_knn.cu:141-155) and the final `(best[0]+best[1]+best[2]) / 3.0f`
 * (simple_knn.cu:183) are restated literally.  The squared distance uses the FMA contraction nvcc applies to
 * `d.x*d.x + d.y*d.y + d.z*d.z`: fma(d.z,d.z, fma(d.x,d.x, d.y*d.y)) (read off the SASS of the reference kernel built with nvcc 12.9).
 * Pinned on the GPU box against the reference itself compiled into oracle/_ref/libsimple_knn_ref.so
 * (tests/test_gpu_knn.py). */
#include <float.h>
#include <math.h>
#include <stdint.h>

void knn_mean_dist2_ref(int32_t P, const float* pts, float* out) {
#pragma omp parallel for schedule(static)
  for (int32_t i = 0; i < P; i++) {
    const float qx = pts[3 * i], qy = pts[3 * i + 1], qz = pts[3 * i + 2];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    for (int32_t j = 0; j < P; j++) {
      if (j == i) continue;
      const float dx = pts[3 * j] - qx, dy = pts[3 * j + 1] - qy, dz = pts[3 * j + 2] - qz;
      float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
      for (int k = 0; k < 3; k++)
        if (best[k] > dist) { const float t = best[k]; best[k] = dist; dist = t; }
    }
    out[i] = (best[0] + best[1] + best[2]) / 3.0f;
  }
}
