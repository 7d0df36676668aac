// ORACLE / TEST INFRASTRUCTURE ONLY.  extern "C" shim around the UNMODIFIED reference
// `SimpleKNN::knn` (submodules/simple-knn/simple_knn.cu:185-221, declared in simple_knn.h), compiled from the
// sources where they lie under /root/reference by oracle/Makefile (target _ref/libsimple_knn_ref.so).
// Used by tests/test_gpu_knn.py as the bit-exact parity target for d3g_knn_mean_dist2; never by the product.
#include <cuda_runtime.h>
#include "simple_knn.h"
extern "C" int ref_simple_knn(int P, float* points_dev, float* mean_dists_dev) {
  SimpleKNN::knn(P, (float3*)points_dev, mean_dists_dev);
  return (int)cudaDeviceSynchronize();
}
