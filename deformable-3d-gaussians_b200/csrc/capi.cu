// C-ABI plumbing shared by every translation unit of libd3g.so: error text, version.
#include <stdarg.h>

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
}  // namespace d3g

extern "C" {
const char* d3g_last_error(void) { return d3g::g_err; }
int d3g_version(void) { return 1; }
}
