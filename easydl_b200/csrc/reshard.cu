// A6 elastic row re-sharding: export the materialised (updated) rows of a shard in ascending
// slab-row order, import records into a shard laid out for a different shard count.
//
// SPEC: export = (keys ascending, rows, slots); import on shard s' keeps
// records with id mod S' == s'.  Pure data movement, bit-exact on every float.  Because the
// row initialiser is a pure function of (seed, table, id), rows that never received an
// update are not exported: the importing shard re-creates them.
//
// Ascending slab row == table-major, ascending id within a table (id = local*S + shard is
// monotonic in the local row), so the compacted stream is already in the order the
// checkpoint format and the all-to-all-v re-shard need.
#include <algorithm>

#include "edl_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWordsPerThread = 8;
constexpr int kTileWords = kThreads * kWordsPerThread;  // 2048 words = 65536 rows per CTA

__device__ __forceinline__ unsigned load_word(const uint32_t* touched, int64_t w, int64_t n_words,
                                              int64_t n_rows) {
  if (w >= n_words) return 0u;
  unsigned v = touched[w];
  int64_t rem = n_rows - w * 32;
  if (rem < 32) v &= (1u << rem) - 1u;
  return v;
}

__global__ void __launch_bounds__(kThreads)
k_popc_tiles(const uint32_t* __restrict__ touched, int64_t n_rows, uint32_t* __restrict__ tile_counts) {
  __shared__ unsigned warp_sum[kThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int64_t n_words = (n_rows + 31) / 32;
  const int64_t base = (int64_t)blockIdx.x * kTileWords + (int64_t)tid * kWordsPerThread;
  unsigned c = 0;
#pragma unroll
  for (int k = 0; k < kWordsPerThread; ++k) c += __popc(load_word(touched, base + k, n_words, n_rows));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) warp_sum[w] = c;
  __syncthreads();
  if (tid == 0) {
    unsigned s = 0;
    for (int i = 0; i < kThreads / 32; ++i) s += warp_sum[i];
    tile_counts[blockIdx.x] = s;
  }
}

__global__ void __launch_bounds__(1024)
k_scan_counts(uint32_t* __restrict__ tile_counts, int n_tiles, int32_t* __restrict__ n_out) {
  __shared__ unsigned warp_sum[32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int per = (n_tiles + 1023) / 1024;
  const int b = tid * per, e = min(b + per, n_tiles);
  unsigned s = 0;
  for (int i = b; i < e; ++i) s += tile_counts[i];
  unsigned inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) warp_sum[w] = inc;
  __syncthreads();
  unsigned wbase = 0, total = 0;
  for (int i = 0; i < 32; ++i) {
    unsigned v = warp_sum[i];
    if (i < w) wbase += v;
    total += v;
  }
  unsigned run = wbase + inc - s;
  for (int i = b; i < e; ++i) {
    unsigned v = tile_counts[i];
    tile_counts[i] = run;
    run += v;
  }
  if (tid == 0) *n_out = (int32_t)total;
}

// writes the slab row of every set bit, ascending, into rows_idx (the caller's keys_out)
__global__ void __launch_bounds__(kThreads)
k_enumerate(const uint32_t* __restrict__ touched, int64_t n_rows, const uint32_t* __restrict__ tile_base,
            int64_t capacity, int64_t* __restrict__ rows_idx) {
  __shared__ unsigned warp_sum[kThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int64_t n_words = (n_rows + 31) / 32;
  const int64_t base = (int64_t)blockIdx.x * kTileWords + (int64_t)tid * kWordsPerThread;
  unsigned word[kWordsPerThread];
  unsigned c = 0;
#pragma unroll
  for (int k = 0; k < kWordsPerThread; ++k) {
    word[k] = load_word(touched, base + k, n_words, n_rows);
    c += __popc(word[k]);
  }
  unsigned inc = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) warp_sum[w] = inc;
  __syncthreads();
  unsigned wbase = 0;
  for (int i = 0; i < kThreads / 32; ++i)
    if (i < w) wbase += warp_sum[i];
  int64_t out = (int64_t)tile_base[blockIdx.x] + wbase + inc - c;
#pragma unroll
  for (int k = 0; k < kWordsPerThread; ++k) {
    unsigned v = word[k];
    while (v) {
      int b = __ffs(v) - 1;
      v &= v - 1;
      if (out < capacity) rows_idx[out] = (base + k) * 32 + b;
      ++out;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
k_copy_rows(const float4* __restrict__ slab, const float4* __restrict__ s0, const float4* __restrict__ s1,
            int D4, int RS4, const int64_t* __restrict__ rows_idx, const int32_t* __restrict__ n_dev,
            int64_t capacity, float4* __restrict__ rows_out, float4* __restrict__ s0_out,
            float4* __restrict__ s1_out) {
  int64_t n = *n_dev;
  if (n > capacity) n = capacity;
  const int64_t total = n * D4;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * kThreads) {
    int64_t i = e / D4;
    int c = (int)(e - i * D4);
    int64_t src = rows_idx[i] * RS4 + c;
    rows_out[e] = slab[src];
    if (s0_out) s0_out[e] = s0[src];
    if (s1_out) s1_out[e] = s1[src];
  }
}

// slab row -> fused key, in place
__global__ void __launch_bounds__(kThreads)
k_rows_to_keys(int64_t* __restrict__ keys, const int32_t* __restrict__ n_dev, int64_t capacity,
               const int64_t* __restrict__ layout) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  int64_t n = *n_dev;
  if (n > capacity) n = capacity;
  if (L.T == 0) return;  // S == 1 without layout: slab row == key
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
    long long row = keys[i];
    int t = edl_find_table(L.lb, L.T, row);
    long long id = (row - L.lb[t]) * (long long)L.S + L.shard;
    keys[i] = L.kb[t] + id;
  }
}

__global__ void __launch_bounds__(kThreads)
k_import(float4* __restrict__ slab, float4* __restrict__ s0, float4* __restrict__ s1, int D4, int RS4,
         uint32_t* __restrict__ touched, const int64_t* __restrict__ layout,
         const int64_t* __restrict__ keys, const float4* __restrict__ rows,
         const float4* __restrict__ s0_in, const float4* __restrict__ s1_in, int64_t n) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  const int64_t total = n * D4;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * kThreads) {
    int64_t i = e / D4;
    int c = (int)(e - i * D4);
    int t; unsigned long long id;
    edl_split_key(&L, keys[i], &t, &id);
    long long row;
    if (L.T == 0) {
      row = (long long)id;
    } else {
      if (id % L.S != L.shard) continue;  // another shard's record
      row = L.lb[t] + (long long)(id / L.S);
    }
    int64_t dst = row * RS4 + c;
    slab[dst] = rows[e];
    if (s0_in) s0[dst] = s0_in[e];
    if (s1_in) s1[dst] = s1_in[e];
    if (c == 0 && touched) atomicOr(&touched[row >> 5], 1u << (row & 31));
  }
}

int n_tiles_for(int64_t n_rows) { return (int)edl_cdiv(edl_cdiv(n_rows, 32), kTileWords); }

int grid_for(int64_t elems) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(edl_cdiv(elems, kThreads), (int64_t)EDL_SM_COUNT * 16));
}

}  // namespace

extern "C" size_t edl_export_workspace_bytes(int64_t n_rows) {
  if (n_rows < 0) return 0;
  return edl_align_up((size_t)(n_tiles_for(n_rows) + 1) * 4, 256);
}

extern "C" int edl_export_count(const uint32_t* touched, int64_t n_rows, int32_t* n_out,
                                void* workspace, size_t workspace_bytes, edl_stream_t stream) {
  EDL_REQUIRE(n_rows >= 0 && n_rows < ((int64_t)1 << 40), EDL_EINVAL, "edl_export_count: n_rows out of range");
  EDL_REQUIRE(n_out, EDL_EINVAL, "edl_export_count: n_out is null");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows == 0) {
    EDL_CUDA(cudaMemsetAsync(n_out, 0, sizeof(int32_t), st));
    return EDL_OK;
  }
  EDL_REQUIRE(touched, EDL_EINVAL, "edl_export_count: touched is null");
  EDL_REQUIRE(workspace && workspace_bytes >= edl_export_workspace_bytes(n_rows), EDL_EWORKSPACE,
              "edl_export_count: workspace %zu < required %zu", workspace_bytes,
              edl_export_workspace_bytes(n_rows));
  const int tiles = n_tiles_for(n_rows);
  uint32_t* tc = (uint32_t*)workspace;
  k_popc_tiles<<<tiles, kThreads, 0, st>>>(touched, n_rows, tc);
  k_scan_counts<<<1, 1024, 0, st>>>(tc, tiles, n_out);
  EDL_LAUNCH_CHECK("export count kernels");
  return EDL_OK;
}

extern "C" int edl_export_shard(const float* slab, const float* slot0, const float* slot1, int dim,
                                int64_t row_stride, const uint32_t* touched, int64_t n_rows, const int64_t* layout,
                                int64_t capacity, int64_t* keys_out, float* rows_out,
                                float* slot0_out, float* slot1_out, int32_t* n_out, void* workspace,
                                size_t workspace_bytes, edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_export_shard: dim=%d must be a positive multiple of 4", dim);
  EDL_REQUIRE(capacity >= 0, EDL_EINVAL, "edl_export_shard: capacity negative");
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_export_shard", dim, row_stride, &RS4); if (rc0) return rc0; }
  int rc = edl_export_count(touched, n_rows, n_out, workspace, workspace_bytes, stream);
  if (rc) return rc;
  if (n_rows == 0 || capacity == 0) return EDL_OK;
  EDL_REQUIRE(slab && keys_out && rows_out, EDL_EINVAL, "edl_export_shard: null pointer");
  EDL_REQUIRE(!slot0_out || slot0, EDL_EINVAL, "edl_export_shard: slot0_out without slot0");
  EDL_REQUIRE(!slot1_out || slot1, EDL_EINVAL, "edl_export_shard: slot1_out without slot1");
  cudaStream_t st = (cudaStream_t)stream;
  const int tiles = n_tiles_for(n_rows);
  const int D4 = dim / 4;
  k_enumerate<<<tiles, kThreads, 0, st>>>(touched, n_rows, (const uint32_t*)workspace, capacity, keys_out);
  k_copy_rows<<<grid_for(capacity * D4), kThreads, 0, st>>>(
      (const float4*)slab, (const float4*)slot0, (const float4*)slot1, D4, RS4, keys_out, n_out, capacity,
      (float4*)rows_out, (float4*)slot0_out, (float4*)slot1_out);
  k_rows_to_keys<<<grid_for(capacity), kThreads, 0, st>>>(keys_out, n_out, capacity, layout);
  EDL_LAUNCH_CHECK("export kernels");
  return EDL_OK;
}

extern "C" int edl_import_shard(float* slab, float* slot0, float* slot1, int dim, int64_t row_stride, uint32_t* touched,
                                const int64_t* layout, const int64_t* keys, const float* rows,
                                const float* slot0_in, const float* slot1_in, int64_t n,
                                edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_import_shard: dim=%d must be a positive multiple of 4", dim);
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_import_shard: n negative");
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_import_shard", dim, row_stride, &RS4); if (rc0) return rc0; }
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(slab && keys && rows, EDL_EINVAL, "edl_import_shard: null pointer");
  EDL_REQUIRE(!slot0_in || slot0, EDL_EINVAL, "edl_import_shard: slot0_in without slot0");
  EDL_REQUIRE(!slot1_in || slot1, EDL_EINVAL, "edl_import_shard: slot1_in without slot1");
  const int D4 = dim / 4;
  k_import<<<grid_for(n * D4), kThreads, 0, (cudaStream_t)stream>>>(
      (float4*)slab, (float4*)slot0, (float4*)slot1, D4, RS4, touched, layout, keys, (const float4*)rows,
      (const float4*)slot0_in, (const float4*)slot1_in, n);
  EDL_LAUNCH_CHECK("k_import");
  return EDL_OK;
}
