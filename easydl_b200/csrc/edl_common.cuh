// Shared device/host helpers for libedlsparse (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/edlsparse.h"

#define EDL_SM_COUNT 148          // B200: 2 dies x 74 SMs
#define EDL_MAX_TABLES 64         // layout decode stages key_base/local_base in shared memory
// Bit 62 of a fused key on the wire marks "this sender has exactly one gradient row for the id" (slot
// exchange, narrow gradient wire).  Lookups by key ignore it; edl_slot_compact strips it.
#define EDL_KEY_FLAG_SINGLE (1ll << 62)

void edl_set_error(const char* fmt, ...);

// Launch-shape switches (api.cu): defaults, overridable by environment variable at the first read and
// by edl_tuning_set() at any time.
enum EdlTune {
  EDL_TUNE_LONG_THR = 0,   // ids with more gradient rows than this go through the chunk kernels
  EDL_TUNE_SEG_FORK,       // 1: hot-id kernels on the helper stream beside the main reduce kernel
  EDL_TUNE_DIRECT_BATCH,   // k_seg_direct: ids per lane group (1 or 2)
  EDL_TUNE_SEG_IMPL,       // 0: k_seg_direct (register path), 1: k_seg_stream (cp.async ring), 2: by gradient row type and width
  EDL_TUNE_STREAM_CTAS,    // k_seg_stream: persistent CTAs per SM
  EDL_TUNE_GATHER_IMPL,    // 0: k_gather, 1: k_gather_pf (row index of the next iteration prefetched)
  EDL_TUNE_CLASSIFY_SIDE,  // forked update: 1 = k_seg_classify on the helper stream too
  EDL_TUNE_STREAM_STAGES,  // k_seg_stream: ring slots in use (2..4)
  EDL_TUNE_DEDUP_IMPL,     // field-major dedup: 2 = one cluster sorts one field in shared memory and writes the
                           // unique outputs, 1 = that sort + separate unique kernels, 0 = global passes
  EDL_TUNE_DEDUP_SHAPE,    // k_cluster_sort: 0 = 4 CTAs x 1024 threads per field, 1 = 8 CTAs x 512 threads
  EDL_TUNE_COUNT
};
int edl_tune(int id);

#define EDL_REQUIRE(cond, code, ...)        \
  do {                                      \
    if (!(cond)) {                          \
      edl_set_error(__VA_ARGS__);           \
      return (code);                        \
    }                                       \
  } while (0)

#define EDL_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      edl_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return EDL_ECUDA;                                                             \
    }                                                                               \
  } while (0)

#define EDL_LAUNCH_CHECK(name)                                                      \
  do {                                                                              \
    cudaError_t e__ = cudaGetLastError();                                           \
    if (e__ != cudaSuccess) {                                                       \
      edl_set_error("launch of %s failed: %s", name, cudaGetErrorString(e__));      \
      return EDL_ECUDA;                                                             \
    }                                                                               \
  } while (0)

// Row stride of a table / slot array in floats: 0 = dense rows (stride == dim).  Parameter and
// optimizer-slot rows of one id may be stored side by side in one record ([p | slot0 | slot1], stride
// (1 + n_slots) * dim, slot pointers offset by dim): one DRAM page and one TLB entry per id instead of
// one per array.  -> stride in float4 units.
static inline int edl_row_stride4(const char* who, int dim, int64_t row_stride, int* out) {
  if (row_stride == 0) row_stride = dim;
  EDL_REQUIRE(row_stride >= dim && row_stride % 4 == 0 && row_stride < (1 << 20), EDL_EINVAL,
              "%s: row_stride=%lld must be 0 or a multiple of 4 that is >= dim=%d", who, (long long)row_stride, dim);
  *out = (int)(row_stride / 4);
  return EDL_OK;
}

static inline size_t edl_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int64_t edl_cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------
// Layout decode.  layout = [T, S, shard, 0, key_base[T+1], local_base[T+1]] (device int64).
// Every kernel that maps fused keys to slab rows copies the two small base arrays into
// shared memory once per CTA.
struct EdlLayoutSmem {
  int T;
  unsigned S;
  unsigned shard;
  long long kb[EDL_MAX_TABLES + 1];
  long long lb[EDL_MAX_TABLES + 1];
};

__device__ __forceinline__ void edl_layout_load(EdlLayoutSmem* L, const int64_t* layout) {
  // caller must __syncthreads() afterwards
  if (layout == nullptr) {
    if (threadIdx.x == 0) { L->T = 0; L->S = 1; L->shard = 0; }
    return;
  }
  int T = (int)layout[0];
  if (threadIdx.x == 0) {
    L->T = T;
    L->S = (unsigned)layout[1];
    L->shard = (unsigned)layout[2];
  }
  for (int i = threadIdx.x; i <= T; i += blockDim.x) {
    L->kb[i] = layout[4 + i];
    L->lb[i] = layout[4 + (T + 1) + i];
  }
}

// largest t with kb[t] <= key  (T <= 64: branch-free binary search over shared memory)
__device__ __forceinline__ int edl_find_table(const long long* base, int T, long long key) {
  int lo = 0, hi = T;  // invariant: base[lo] <= key < base[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (base[mid] <= key) lo = mid; else hi = mid;
  }
  return lo;
}

// fused key -> (table, per-table id)
__device__ __forceinline__ void edl_split_key(const EdlLayoutSmem* L, long long key, int* t,
                                              unsigned long long* id) {
  if (L->T == 0) { *t = 0; *id = (unsigned long long)key; return; }
  int tt = edl_find_table(L->kb, L->T, key);
  *t = tt;
  *id = (unsigned long long)(key - L->kb[tt]);
}

// fused key -> slab row of this shard (does not check ownership)
__device__ __forceinline__ long long edl_key_to_row(const EdlLayoutSmem* L, long long key) {
  if (L->T == 0) return key;
  int t; unsigned long long id;
  edl_split_key(L, key, &t, &id);
  return L->lb[t] + (long long)(id / L->S);
}

__device__ __forceinline__ unsigned edl_lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// streaming 128-bit store (evict-first: output buffers are written once and not re-read
// by this kernel; keep L2 for table rows and id tiles)
__device__ __forceinline__ void edl_st_cs(float4* p, float4 v) { __stcs(p, v); }
__device__ __forceinline__ float4 edl_ld_cs(const float4* p) { return __ldcs(p); }
