// Library-wide pieces of the C ABI: version, the thread-local error string and the tuning knobs.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "edl_common.cuh"

namespace {
thread_local char g_err[512] = "";

// One table for every launch-shape switch of the library.  A knob starts at its default, or at the
// value of its environment variable when that is set at the first read, and can be changed at any
// time with edl_tuning_set (atomics: a concurrent launch sees the old or the new value).
struct Knob { const char* name; const char* env; int def, lo, hi; };
const Knob kKnobs[EDL_TUNE_COUNT] = {
    /* EDL_TUNE_LONG_THR      */ {"long_thr", "EDL_LONG_THR", 16, 2, 1 << 20},
    /* EDL_TUNE_SEG_FORK      */ {"seg_fork", "EDL_SEG_FORK", 1, 0, 1},
    /* EDL_TUNE_DIRECT_BATCH  */ {"direct_batch", "EDL_DIRECT_BATCH", 2, 1, 2},
    /* EDL_TUNE_SEG_IMPL      */ {"seg_impl", "EDL_SEG_IMPL", 2, 0, 2},
    /* EDL_TUNE_STREAM_CTAS   */ {"stream_ctas", "EDL_STREAM_CTAS", 3, 1, 8},
    /* EDL_TUNE_GATHER_IMPL   */ {"gather_impl", "EDL_GATHER_IMPL", 1, 0, 1},
    /* EDL_TUNE_CLASSIFY_SIDE */ {"classify_side", "EDL_CLASSIFY_SIDE", 1, 0, 1},
    /* EDL_TUNE_STREAM_STAGES */ {"stream_stages", "EDL_STREAM_STAGES", 4, 2, 4},
    /* EDL_TUNE_DEDUP_IMPL    */ {"dedup_impl", "EDL_DEDUP_IMPL", 2, 0, 2},
    /* EDL_TUNE_DEDUP_SHAPE   */ {"dedup_shape", "EDL_DEDUP_SHAPE", 0, 0, 1},
};
std::atomic<int> g_knob[EDL_TUNE_COUNT];
std::once_flag g_knob_once;

void knobs_init() {
  for (int i = 0; i < EDL_TUNE_COUNT; ++i) {
    int v = kKnobs[i].def;
    const char* e = getenv(kKnobs[i].env);
    if (e && *e) v = atoi(e);
    if (v < kKnobs[i].lo) v = kKnobs[i].lo;
    if (v > kKnobs[i].hi) v = kKnobs[i].hi;
    g_knob[i].store(v, std::memory_order_relaxed);
  }
}
int knob_index(const char* name) {
  if (!name) return -1;
  for (int i = 0; i < EDL_TUNE_COUNT; ++i)
    if (strcmp(name, kKnobs[i].name) == 0) return i;
  return -1;
}
}  // namespace

void edl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int edl_tune(int id) {
  std::call_once(g_knob_once, knobs_init);
  return g_knob[id].load(std::memory_order_relaxed);
}

extern "C" int edl_version(void) { return 200; }  // 0.2.0

extern "C" const char* edl_last_error_string(void) { return g_err; }

extern "C" int edl_tuning_set(const char* name, int value) {
  std::call_once(g_knob_once, knobs_init);
  const int i = knob_index(name);
  EDL_REQUIRE(i >= 0, EDL_EINVAL, "edl_tuning_set: unknown knob '%s'", name ? name : "(null)");
  EDL_REQUIRE(value >= kKnobs[i].lo && value <= kKnobs[i].hi, EDL_EINVAL,
              "edl_tuning_set: %s=%d outside %d..%d", name, value, kKnobs[i].lo, kKnobs[i].hi);
  g_knob[i].store(value, std::memory_order_relaxed);
  return EDL_OK;
}

extern "C" int edl_tuning_get(const char* name, int* value) {
  std::call_once(g_knob_once, knobs_init);
  const int i = knob_index(name);
  EDL_REQUIRE(i >= 0 && value, EDL_EINVAL, "edl_tuning_get: unknown knob '%s' or null output", name ? name : "(null)");
  *value = g_knob[i].load(std::memory_order_relaxed);
  return EDL_OK;
}
