// C-ABI plumbing shared by every translation unit of libd3g.so: error text, version.
#include <stdarg.h>

#include <vector>

#include "common.cuh"

namespace d3g {
thread_local char g_err[512] = "";
int set_err(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

bool g_prof_on = false;
struct ProfRec { int id; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_pool;
static cudaEvent_t prof_event() {
  if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
void prof_begin(int id, cudaStream_t stream) {
  ProfRec r; r.id = id; r.a = prof_event(); r.b = prof_event();
  cudaEventRecord(r.a, stream);
  g_prof_recs.push_back(r);
}
void prof_end(int id, cudaStream_t stream) {
  for (size_t i = g_prof_recs.size(); i-- > 0;)
    if (g_prof_recs[i].id == id) { cudaEventRecord(g_prof_recs[i].b, stream); return; }
}
}  // namespace d3g

static const char* kProfNames[d3g::PROF_COUNT] = {
    "preprocess_fwd", "scan", "duplicate_with_keys", "pad_keys", "radix_sort", "identify_tile_ranges", "render_fwd",
    "render_bwd", "preprocess_bwd", "mlp_tfeat", "mlp_fwd", "mlp_dgrad", "mlp_wgrad", "mlp_finish", "mlp_pack_weights", "image_loss_fwd", "image_loss_bwd", "adam_step", "mlp_input_bwd", "densification_stats", "tile_sort", "tile_count", "tile_scan", "tile_scatter", "mlp_amax"};

extern "C" {
int d3g_prof_enable(int32_t on) { d3g::g_prof_on = on != 0; return 0; }
int d3g_prof_count(void) { return d3g::PROF_COUNT; }
const char* d3g_prof_name(int32_t id) { return (id >= 0 && id < d3g::PROF_COUNT) ? kProfNames[id] : ""; }
int d3g_prof_collect(double* total_ms, int32_t* launches, int32_t n) {
  using namespace d3g;
  if (!total_ms || !launches || n < PROF_COUNT) return set_err(D3G_ERR_INVALID, "bad arguments");
  for (int i = 0; i < n; i++) { total_ms[i] = 0.0; launches[i] = 0; }
  D3G_CUDA_TRY(cudaDeviceSynchronize());
  for (auto& r : g_prof_recs) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { total_ms[r.id] += ms; launches[r.id]++; }
    g_prof_pool.push_back(r.a);
    g_prof_pool.push_back(r.b);
  }
  g_prof_recs.clear();
  return 0;
}
const char* d3g_last_error(void) { return d3g::g_err; }
int d3g_version(void) { return 2; }
}
