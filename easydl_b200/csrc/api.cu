// Library-wide pieces of the C ABI: version and the thread-local error string.
#include <stdarg.h>

#include "edl_common.cuh"

namespace {
thread_local char g_err[512] = "";
}

void edl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" int edl_version(void) { return 100; }  // 0.1.0

extern "C" const char* edl_last_error_string(void) { return g_err; }
