// placeholder until the backward kernels land (next commit)
#include "mlp_common.cuh"
using namespace d3g;
extern "C" int d3g_mlp_backward(const d3g_mlp_config*, const void*, const d3g_mlp_params*, int64_t, const float*,
                                const float*, int64_t, void*, const float*, const float*, const float*,
                                const d3g_mlp_grads*, float*, float*, void*) {
  return set_err(D3G_ERR_UNSUPPORTED, "d3g_mlp_backward: not built yet");
}
