// Lazily materialised rows (SURVEY.md 7 hard part 1, 8(f).3): BASELINE.json configs[2] is 26 tables
// x 40M rows x dim 128 with Adam = 1.6 TB dense, which 8 x 180 GB cannot hold.  Because the row
// initialiser is a pure function of (seed, table, id), a shard only has to store the rows that have
// received an update: an open-addressing hash index (fused key -> slab slot, linear probing, keys
// claimed with a 64-bit CAS) in front of a slab that grows by an atomic counter.  Lookups of rows
// that were never updated compute the initial value on the fly and store nothing.  This mirrors the
// lineage's EmbeddingTable.get lazy initialisation, minus the Python dict.
#include <algorithm>

#include "edl_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr long long kEmpty = -1;

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  unsigned long long z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ float init_value(unsigned long long salt, unsigned long long idx) {
  unsigned long long h = mix64(salt ^ idx);
  float x = __fmul_rn((float)(unsigned)(h >> 40), 5.9604644775390625e-08f);
  return __fsub_rn(__fmul_rn(x, 0.1f), 0.05f);
}

__device__ __forceinline__ float4 init_row4(const EdlLayoutSmem* L, long long key, int D4, int c4,
                                            unsigned long long seed) {
  int t; unsigned long long id;
  edl_split_key(L, key, &t, &id);
  const unsigned long long salt = seed ^ ((unsigned long long)t << 56);
  const unsigned long long idx = id * (unsigned long long)(D4 * 4) + (unsigned long long)c4 * 4ull;
  return make_float4(init_value(salt, idx), init_value(salt, idx + 1), init_value(salt, idx + 2),
                     init_value(salt, idx + 3));
}

// read-only probe: slab slot of `key` or -1
__device__ __forceinline__ long long probe(const long long* __restrict__ idx_keys,
                                           const int* __restrict__ idx_slot, unsigned long long mask,
                                           long long key) {
  unsigned long long h = mix64((unsigned long long)key) & mask;
  for (unsigned long long n = 0; n <= mask; ++n) {
    const long long k = idx_keys[h];
    if (k == key) return idx_slot[h];
    if (k == kEmpty) return -1;
    h = (h + 1) & mask;
  }
  return -1;
}

// A3 on a lazy shard: stored row, or the initial value computed on the fly
template <int kD4>
__global__ void __launch_bounds__(kThreads, 6)
k_lazy_gather(const long long* __restrict__ idx_keys, const int* __restrict__ idx_slot,
              unsigned long long mask, const float4* __restrict__ slab, int D4_rt, int RS4,
              const int64_t* __restrict__ keys, int64_t n, const int32_t* n_dev,
              const int64_t* __restrict__ layout, unsigned long long seed, float4* __restrict__ out) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  const int D4 = kD4 > 0 ? kD4 : D4_rt;
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  const int64_t total = ne * D4;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
    const int64_t i = e / D4;
    const int c = (int)(e - i * D4);
    const long long key = keys[i] & ~EDL_KEY_FLAG_SINGLE;
    const long long slot = probe(idx_keys, idx_slot, mask, key);
    out[e] = slot >= 0 ? slab[slot * RS4 + c] : init_row4(&L, key, D4, c, seed);
  }
}

// insert-if-absent for UNIQUE keys (the output of A1): slots_out[i] = slab slot of keys[i];
// a new key claims the next slot and is flagged for initialisation.
__global__ void __launch_bounds__(kThreads)
k_lazy_insert(long long* __restrict__ idx_keys, int* __restrict__ idx_slot, unsigned long long mask,
              const int64_t* __restrict__ keys, int64_t n, const int32_t* n_dev, int* __restrict__ count,
              int max_rows, int64_t* __restrict__ slots_out, int* __restrict__ err) {
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < ne; i += (int64_t)gridDim.x * kThreads) {
    const long long key = keys[i];
    unsigned long long h = mix64((unsigned long long)key) & mask;
    long long slot = -1;
    for (unsigned long long probes = 0; probes <= mask; ++probes) {
      const long long k = idx_keys[h];
      if (k == key) { slot = idx_slot[h]; break; }
      if (k == kEmpty) {
        const long long old = (long long)atomicCAS((unsigned long long*)&idx_keys[h], (unsigned long long)kEmpty,
                                                   (unsigned long long)key);
        if (old == kEmpty) {
          const int s = atomicAdd(count, 1);
          if (s >= max_rows) {
            // slab full: the key stays in the index but points at the spare row (slab row max_rows), the
            // count is put back, the error flag is raised -- later probes of this key read a defined slot
            atomicSub(count, 1);
            idx_slot[h] = max_rows;
            atomicExch(err, 1);
            slot = max_rows;
            break;
          }
          idx_slot[h] = s;
          slot = (long long)s | (1ll << 62);   // bit 62: freshly claimed, needs initialisation
          break;
        }
        if (old == key) { slot = idx_slot[h]; break; }   // cannot happen for unique keys; harmless
      }
      h = (h + 1) & mask;
    }
    if (slot == -1) atomicExch(err, 1);
    slots_out[i] = slot;
  }
}

// initialise the rows claimed by k_lazy_insert and strip the marker bit
__global__ void __launch_bounds__(kThreads)
k_lazy_init_new(float4* __restrict__ slab, float4* __restrict__ slot0, float4* __restrict__ slot1, int D4, int RS4,
                const int64_t* __restrict__ keys, int64_t* __restrict__ slots, int64_t n, const int32_t* n_dev,
                const int64_t* __restrict__ layout, unsigned long long seed, float init0, float init1) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  const int64_t total = ne * D4;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (int64_t)gridDim.x * kThreads) {
    const int64_t i = e / D4;
    const int c = (int)(e - i * D4);
    const long long s = slots[i];
    if (s >= 0 && (s & (1ll << 62))) {
      const long long row = s & ~(1ll << 62);
      slab[row * RS4 + c] = init_row4(&L, keys[i], D4, c, seed);
      if (slot0) slot0[row * RS4 + c] = make_float4(init0, init0, init0, init0);
      if (slot1) slot1[row * RS4 + c] = make_float4(init1, init1, init1, init1);
    }
  }
}

// strip the marker bit; keys that found no room (overflow, reported through the err flag) are
// pointed at the spare row `max_rows` so that the update that follows stays inside the slab
__global__ void __launch_bounds__(kThreads)
k_lazy_strip(int64_t* __restrict__ slots, int64_t n, const int32_t* n_dev, long long spare_row) {
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < ne; i += (int64_t)gridDim.x * kThreads) {
    const long long s = slots[i];
    slots[i] = s >= 0 ? (s & ~(1ll << 62)) : spare_row;
  }
}

// all (key, slot) pairs of the index, unordered: the caller sorts by key for export
__global__ void __launch_bounds__(kThreads)
k_lazy_dump(const long long* __restrict__ idx_keys, const int* __restrict__ idx_slot, unsigned long long cap,
            int64_t* __restrict__ keys_out, int64_t* __restrict__ slots_out) {
  for (unsigned long long h = (unsigned long long)blockIdx.x * kThreads + threadIdx.x; h < cap;
       h += (unsigned long long)gridDim.x * kThreads) {
    const long long k = idx_keys[h];
    if (k != kEmpty) {
      const int s = idx_slot[h];
      keys_out[s] = k;          // slots are dense 0..count-1: slot order is a valid enumeration
      slots_out[s] = s;
    }
  }
}

int grid_for(int64_t elems) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(edl_cdiv(elems, kThreads), (int64_t)EDL_SM_COUNT * 16));
}

bool pow2(int64_t x) { return x > 0 && (x & (x - 1)) == 0; }

}  // namespace

extern "C" size_t edl_lazy_index_bytes(int64_t index_capacity) {
  if (!pow2(index_capacity)) return 0;
  return (size_t)index_capacity * 12 + 256;   // int64 keys | int32 slots | [count, err]
}

// index layout inside the caller's buffer: keys[cap] i64, slots[cap] i32, then int32 count, err
struct LazyIdx { long long* keys; int* slots; int* count; int* err; };
static LazyIdx carve_idx(void* index, int64_t cap) {
  LazyIdx r; char* p = (char*)index;
  r.keys = (long long*)p; r.slots = (int*)(p + (size_t)cap * 8);
  r.count = (int*)(p + (size_t)cap * 12); r.err = r.count + 1;
  return r;
}

extern "C" int edl_lazy_clear(void* index, int64_t index_capacity, edl_stream_t stream) {
  EDL_REQUIRE(index && pow2(index_capacity), EDL_EINVAL, "edl_lazy_clear: index null or capacity not a power of two");
  LazyIdx I = carve_idx(index, index_capacity);
  EDL_CUDA(cudaMemsetAsync(I.keys, 0xFF, (size_t)index_capacity * 8, (cudaStream_t)stream));
  EDL_CUDA(cudaMemsetAsync(I.slots, 0, (size_t)index_capacity * 4, (cudaStream_t)stream));
  EDL_CUDA(cudaMemsetAsync(I.count, 0, 8, (cudaStream_t)stream));
  return EDL_OK;
}

extern "C" int edl_lazy_gather(const void* index, int64_t index_capacity, const float* slab, int dim,
                               int64_t row_stride, const int64_t* keys, int64_t n, const int32_t* n_dev, const int64_t* layout,
                               uint64_t seed, float* out, edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_lazy_gather: dim=%d must be a positive multiple of 4", dim);
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_lazy_gather: n negative");
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_lazy_gather", dim, row_stride, &RS4); if (rc0) return rc0; }
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(index && pow2(index_capacity) && slab && keys && out && layout, EDL_EINVAL, "edl_lazy_gather: null pointer");
  LazyIdx I = carve_idx((void*)index, index_capacity);
  const int D4 = dim / 4;
  const int g = grid_for(n * D4);
  cudaStream_t st = (cudaStream_t)stream;
#define EDL_CASE(W) case W: k_lazy_gather<W><<<g, kThreads, 0, st>>>(I.keys, I.slots, (unsigned long long)index_capacity - 1, (const float4*)slab, D4, RS4, keys, n, n_dev, layout, seed, (float4*)out); break;
  switch (D4) {
    EDL_CASE(4) EDL_CASE(8) EDL_CASE(16) EDL_CASE(32) EDL_CASE(64)
    default: k_lazy_gather<0><<<g, kThreads, 0, st>>>(I.keys, I.slots, (unsigned long long)index_capacity - 1, (const float4*)slab, D4, RS4, keys, n, n_dev, layout, seed, (float4*)out);
  }
#undef EDL_CASE
  EDL_LAUNCH_CHECK("k_lazy_gather");
  return EDL_OK;
}

extern "C" int edl_lazy_insert(void* index, int64_t index_capacity, float* slab, float* slot0, float* slot1,
                               int dim, int64_t row_stride, int64_t max_rows, const int64_t* keys, int64_t n, const int32_t* n_dev,
                               const int64_t* layout, uint64_t seed, float slot0_init, float slot1_init,
                               int64_t* slots_out, edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_lazy_insert: dim=%d must be a positive multiple of 4", dim);
  EDL_REQUIRE(n >= 0 && max_rows > 0 && max_rows < 0x7fffffff, EDL_EINVAL, "edl_lazy_insert: bad sizes");
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_lazy_insert", dim, row_stride, &RS4); if (rc0) return rc0; }
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(index && pow2(index_capacity) && slab && keys && slots_out && layout, EDL_EINVAL, "edl_lazy_insert: null pointer");
  EDL_REQUIRE(index_capacity >= 2 * max_rows, EDL_EINVAL, "edl_lazy_insert: index capacity must be >= 2 x max_rows");
  LazyIdx I = carve_idx(index, index_capacity);
  cudaStream_t st = (cudaStream_t)stream;
  const int D4 = dim / 4;
  k_lazy_insert<<<grid_for(n), kThreads, 0, st>>>(I.keys, I.slots, (unsigned long long)index_capacity - 1, keys, n,
                                                  n_dev, I.count, (int)max_rows, slots_out, I.err);
  k_lazy_init_new<<<grid_for(n * D4), kThreads, 0, st>>>((float4*)slab, (float4*)slot0, (float4*)slot1, D4, RS4, keys,
                                                        slots_out, n, n_dev, layout, seed, slot0_init, slot1_init);
  k_lazy_strip<<<grid_for(n), kThreads, 0, st>>>(slots_out, n, n_dev, (long long)max_rows);
  EDL_LAUNCH_CHECK("lazy insert kernels");
  return EDL_OK;
}

// stamps[slots[i]] = value for the rows of one update (i < *n_dev): the "last updated at step" mark that
// an eviction policy reads
namespace {
__global__ void __launch_bounds__(kThreads)
k_lazy_stamp(int32_t* __restrict__ stamps, const int64_t* __restrict__ slots, int64_t n, const int32_t* n_dev,
             int32_t value) {
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < ne; i += (int64_t)gridDim.x * kThreads)
    stamps[slots[i]] = value;
}
}  // namespace

extern "C" int edl_lazy_stamp(int32_t* stamps, const int64_t* slots, int64_t n, const int32_t* n_dev, int32_t value,
                              edl_stream_t stream) {
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_lazy_stamp: n negative");
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(stamps && slots, EDL_EINVAL, "edl_lazy_stamp: null pointer");
  k_lazy_stamp<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(stamps, slots, n, n_dev, value);
  EDL_LAUNCH_CHECK("k_lazy_stamp");
  return EDL_OK;
}

// count (rows materialised) and err (1 = slab or index full) as a device int32[2]
extern "C" const int32_t* edl_lazy_status_ptr(const void* index, int64_t index_capacity) {
  if (!index || !pow2(index_capacity)) return nullptr;
  return (const int32_t*)carve_idx((void*)index, index_capacity).count;
}

extern "C" int edl_lazy_dump(const void* index, int64_t index_capacity, int64_t* keys_out, int64_t* slots_out,
                             edl_stream_t stream) {
  EDL_REQUIRE(index && pow2(index_capacity) && keys_out && slots_out, EDL_EINVAL, "edl_lazy_dump: null pointer");
  LazyIdx I = carve_idx((void*)index, index_capacity);
  k_lazy_dump<<<grid_for(index_capacity), kThreads, 0, (cudaStream_t)stream>>>(I.keys, I.slots,
                                                                               (unsigned long long)index_capacity,
                                                                               keys_out, slots_out);
  EDL_LAUNCH_CHECK("k_lazy_dump");
  return EDL_OK;
}
