// Row movement kernels: deterministic row initialisation, A3 gather (by fused key, by raw
// [B, F] ids, by int32 index = worker-side expand) and the scatter used to fill all-to-all-v
// send buffers.  Pure data movement, bit-exact (SPEC A3).
//
// Mapping: one thread per 128-bit (float4) element of the OUTPUT, flat index e = row*D4 + c,
// so a warp writes 512 contiguous bytes and reads D4-lane-wide contiguous row pieces
// (256 B for dim 64).  Each thread keeps kUnroll independent row loads in flight; with
// 2048 resident threads per SM that is 128 KB of loads outstanding per SM, above the
// ~44 KB/SM that HBM latency x bandwidth needs on B200.
#include <algorithm>

#include "edl_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 2;

// ------------------------------------------------------------------ row init
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  unsigned long long z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ float init_value(unsigned long long salt, unsigned long long idx) {
  unsigned long long h = splitmix64(salt ^ idx);
  float x = __fmul_rn((float)(unsigned)(h >> 40), 5.9604644775390625e-08f);  // k * 2^-24, exact
  return __fsub_rn(__fmul_rn(x, 0.1f), 0.05f);                              // two roundings, no FMA
}

__global__ void __launch_bounds__(kThreads)
k_init_rows(float4* __restrict__ slab, int D4, int RS4, const int64_t* __restrict__ layout, int64_t row_begin,
            int64_t n_rows, unsigned long long seed) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  const int64_t total = n_rows * D4;
  const unsigned long long D = (unsigned long long)D4 * 4ull;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    int64_t row = row_begin + e / D4;
    int c4 = (int)(e % D4);
    int t = edl_find_table(L.lb, L.T, row);
    unsigned long long id = (unsigned long long)(row - L.lb[t]) * L.S + L.shard;
    unsigned long long salt = seed ^ ((unsigned long long)t << 56);
    unsigned long long idx = id * D + (unsigned long long)c4 * 4ull;
    float4 v;
    v.x = init_value(salt, idx + 0);
    v.y = init_value(salt, idx + 1);
    v.z = init_value(salt, idx + 2);
    v.w = init_value(salt, idx + 3);
    slab[row * RS4 + c4] = v;
  }
}

// value fill of `dim`-wide rows that lie `RS4` float4 apart (optimizer slots inside interleaved records)
__global__ void __launch_bounds__(kThreads)
k_fill_rows(float4* __restrict__ dst, int D4, int RS4, int64_t n_rows, float value) {
  const float4 v = make_float4(value, value, value, value);
  const int64_t total = n_rows * D4;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / D4;
    dst[r * RS4 + (e - r * D4)] = v;
  }
}

__global__ void __launch_bounds__(kThreads)
k_fill(float* __restrict__ dst, int64_t n, float value) {
  const int64_t n4 = n >> 2;
  float4 v = make_float4(value, value, value, value);
  float4* d4 = (float4*)dst;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * blockDim.x)
    d4[i] = v;
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) dst[(n4 << 2) + threadIdx.x] = value;
}

// ------------------------------------------------------------------ gather family
struct KeyRowFn {  // fused key -> slab row through the shard layout
  const int64_t* keys;
  __device__ __forceinline__ long long row(const EdlLayoutSmem* L, int64_t i) const {
    return edl_key_to_row(L, keys[i] & ~EDL_KEY_FLAG_SINGLE);
  }
};
struct IdsRowFn {  // raw ids[B, F] on one shard: slab row = id + field_base[f]
  const int64_t* ids;
  const int64_t* field_base;
  int n_fields;
  __device__ __forceinline__ long long row(const EdlLayoutSmem*, int64_t j) const {
    long long r = ids[j];
    if (field_base) r += field_base[(int)(j % n_fields)];
    return r;
  }
};
// The same with j % n_fields as a multiply-high: q = umulhi(j, M), M = floor(2^32 / d) + 1, is
// floor(j / d) whenever j * (M * d - 2^32) < 2^32 (checked on the host); the 64-bit modulo above costs
// ~40 instructions per element, 16 lanes of a row repeating it.
struct IdsRowFnFast {
  const int64_t* ids;
  const int64_t* field_base;
  unsigned magic, nf;
  __device__ __forceinline__ long long row(const EdlLayoutSmem*, int64_t j) const {
    const unsigned ju = (unsigned)j;
    const unsigned f = ju - __umulhi(ju, magic) * nf;
    return ids[j] + field_base[f];
  }
};
struct IndexRowFn {  // int32 index (expand by `inverse`, permute by `perm`)
  const int32_t* index;
  __device__ __forceinline__ long long row(const EdlLayoutSmem*, int64_t j) const { return index[j]; }
};

// kD4 > 0: compile-time row width in float4 (power of two); kD4 == 0: runtime width.
template <int kD4, typename RowFn, bool kUseLayout>
__global__ void __launch_bounds__(kThreads, 8)
k_gather(const float4* __restrict__ src, int D4_rt, int S4, RowFn fn, int64_t n, const int32_t* n_dev,
         const int64_t* __restrict__ layout, float4* __restrict__ out) {
  __shared__ EdlLayoutSmem L;
  if (kUseLayout) {
    edl_layout_load(&L, layout);
    __syncthreads();
  }
  const int D4 = kD4 > 0 ? kD4 : D4_rt;
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  const int64_t total = ne * D4;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t e0 = (int64_t)blockIdx.x * kThreads + threadIdx.x; e0 < total; e0 += stride * kUnroll) {
    long long r[kUnroll];
    int c[kUnroll];
    float4 v[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      int64_t e = e0 + k * stride;
      if (e < total) {
        int64_t i = e / D4;
        c[k] = (int)(e - i * D4);
        r[k] = fn.row(&L, i);
      }
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      int64_t e = e0 + k * stride;
      if (e < total) v[k] = src[r[k] * S4 + c[k]];
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      int64_t e = e0 + k * stride;
      if (e < total) edl_st_cs(out + e, v[k]);
    }
  }
}

// The same gather with the row index of the NEXT iteration fetched while the rows of this one are in
// flight: the id / key / index load (and the table search behind a key) leaves the critical path, so a
// thread's two row loads are outstanding nearly all the time instead of every other memory round trip.
// For power-of-two row widths that divide the CTA the column of a thread is fixed (c = tid mod D4).
template <int kD4, typename RowFn, bool kUseLayout>
__global__ void __launch_bounds__(kThreads, 8)
k_gather_pf(const float4* __restrict__ src, int D4_rt, int S4, RowFn fn, int64_t n, const int32_t* n_dev,
            const int64_t* __restrict__ layout, float4* __restrict__ out) {
  __shared__ EdlLayoutSmem L;
  if (kUseLayout) {
    edl_layout_load(&L, layout);
    __syncthreads();
  }
  const int D4 = kD4 > 0 ? kD4 : D4_rt;
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  const int64_t total = ne * D4;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  int64_t e0 = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  long long r[kUnroll], rn[kUnroll];
#pragma unroll
  for (int k = 0; k < kUnroll; ++k) {
    const int64_t e = e0 + k * stride;
    r[k] = e < total ? fn.row(&L, e / D4) : 0;
  }
  for (; e0 < total; e0 += stride * kUnroll) {
    const int64_t en = e0 + stride * kUnroll;
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      const int64_t e = en + k * stride;
      rn[k] = e < total ? fn.row(&L, e / D4) : 0;
    }
    float4 v[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      const int64_t e = e0 + k * stride;
      if (e < total) v[k] = src[r[k] * S4 + (int)(e % D4)];
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      const int64_t e = e0 + k * stride;
      if (e < total) edl_st_cs(out + e, v[k]);
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) r[k] = rn[k];
  }
}

// expand into a window of fields: item i = (sample b, field f of the window) lands at output row
// b * F + f0 + f, so that several field groups of one [B, F] batch fill one [B * F, dim] array.
template <int kD4>
__global__ void __launch_bounds__(kThreads, 8)
k_expand_window(const float4* __restrict__ src, int D4_rt, const int32_t* __restrict__ index, int64_t n,
                int Fw, int F, int f0, float4* __restrict__ out) {
  const int D4 = kD4 > 0 ? kD4 : D4_rt;
  const int64_t total = n * D4;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t e0 = (int64_t)blockIdx.x * kThreads + threadIdx.x; e0 < total; e0 += stride * kUnroll) {
    unsigned r[kUnroll], o[kUnroll];   // float4 offsets; the entry point checks that they fit 32 bits
    float4 v[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      int64_t e = e0 + k * stride;
      if (e < total) {
        const unsigned i = (unsigned)(e / D4);
        const unsigned c = (unsigned)e - i * (unsigned)D4;
        const unsigned b = i / (unsigned)Fw;
        r[k] = (unsigned)index[i] * (unsigned)D4 + c;
        o[k] = (b * (unsigned)F + (unsigned)f0 + (i - b * (unsigned)Fw)) * (unsigned)D4 + c;
      }
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      int64_t e = e0 + k * stride;
      if (e < total) v[k] = src[r[k]];
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      int64_t e = e0 + k * stride;
      if (e < total) edl_st_cs(out + o[k], v[k]);
    }
  }
}

// scatter form: out[pos[i]] = rows[i]
template <int kD4>
__global__ void __launch_bounds__(kThreads)
k_scatter_rows(const float4* __restrict__ rows, int D4_rt, int OS4, const int32_t* __restrict__ pos, int64_t n,
               const int32_t* n_dev, float4* __restrict__ out) {
  const int D4 = kD4 > 0 ? kD4 : D4_rt;
  int64_t ne = n;
  if (n_dev) { int64_t m = *n_dev; ne = m < n ? m : n; }
  const int64_t total = ne * D4;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * kThreads) {
    int64_t i = e / D4;
    int c = (int)(e - i * D4);
    out[(int64_t)pos[i] * OS4 + c] = edl_ld_cs(rows + e);
  }
}

__global__ void __launch_bounds__(kThreads)
k_compose(const int32_t* __restrict__ index, const int32_t* __restrict__ map, int64_t n,
          int32_t* __restrict__ out) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n;
       j += (int64_t)gridDim.x * blockDim.x)
    out[j] = map[index[j]];
}

int grid_for(int64_t elems, int per_thread) {
  int64_t ctas = edl_cdiv(elems, (int64_t)kThreads * per_thread);
  // whole waves: 148 SMs x 8 resident CTAs of 256 threads
  const int64_t wave = (int64_t)EDL_SM_COUNT * 8;
  if (ctas > wave) ctas = std::min<int64_t>(edl_cdiv(ctas, wave) * wave, wave * 4);
  return (int)std::max<int64_t>(ctas, 1);
}

template <typename RowFn, bool kUseLayout>
int launch_gather(const float* src, int dim, int64_t row_stride, RowFn fn, int64_t n, const int32_t* n_dev,
                  const int64_t* layout, float* out, cudaStream_t st, const char* who) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "%s: dim=%d must be a positive multiple of 4", who, dim);
  int S4 = 0;
  { int rc0 = edl_row_stride4(who, dim, row_stride, &S4); if (rc0) return rc0; }
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "%s: n=%lld negative", who, (long long)n);
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(src && out, EDL_EINVAL, "%s: null pointer", who);
  EDL_REQUIRE((((uintptr_t)src | (uintptr_t)out) & 15) == 0, EDL_EINVAL, "%s: rows must be 16-byte aligned", who);
  const int D4 = dim / 4;
  const int g = grid_for(n * D4, kUnroll);
  const size_t pad = 0;
  const float4* s4 = (const float4*)src;
  float4* o4 = (float4*)out;
  if (edl_tune(EDL_TUNE_GATHER_IMPL) == 1) {
    switch (D4) {
#define EDL_CASE(W) case W: k_gather_pf<W, RowFn, kUseLayout><<<g, kThreads, pad, st>>>(s4, D4, S4, fn, n, n_dev, layout, o4); break;
      EDL_CASE(4) EDL_CASE(8) EDL_CASE(16) EDL_CASE(32) EDL_CASE(64)
#undef EDL_CASE
      default: k_gather_pf<0, RowFn, kUseLayout><<<g, kThreads, pad, st>>>(s4, D4, S4, fn, n, n_dev, layout, o4);
    }
  } else {
    switch (D4) {
#define EDL_CASE(W) case W: k_gather<W, RowFn, kUseLayout><<<g, kThreads, pad, st>>>(s4, D4, S4, fn, n, n_dev, layout, o4); break;
      EDL_CASE(4) EDL_CASE(8) EDL_CASE(16) EDL_CASE(32) EDL_CASE(64)
#undef EDL_CASE
      default: k_gather<0, RowFn, kUseLayout><<<g, kThreads, pad, st>>>(s4, D4, S4, fn, n, n_dev, layout, o4);
    }
  }
  EDL_LAUNCH_CHECK(who);
  return EDL_OK;
}

}  // namespace

extern "C" int edl_init_rows(float* slab, int dim, int64_t row_stride, const int64_t* layout, int64_t row_begin,
                             int64_t n_rows, uint64_t seed, edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_init_rows: dim=%d must be a positive multiple of 4", dim);
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_init_rows", dim, row_stride, &RS4); if (rc0) return rc0; }
  EDL_REQUIRE(layout, EDL_EINVAL, "edl_init_rows: layout is required (tables are located through local_base)");
  EDL_REQUIRE(row_begin >= 0 && n_rows >= 0, EDL_EINVAL, "edl_init_rows: negative range");
  if (n_rows == 0) return EDL_OK;
  EDL_REQUIRE(slab && (((uintptr_t)slab) & 15) == 0, EDL_EINVAL, "edl_init_rows: slab null or misaligned");
  const int D4 = dim / 4;
  const int g = grid_for(n_rows * D4, 1);
  k_init_rows<<<g, kThreads, 0, (cudaStream_t)stream>>>((float4*)slab, D4, RS4, layout, row_begin, n_rows, seed);
  EDL_LAUNCH_CHECK("k_init_rows");
  return EDL_OK;
}

extern "C" int edl_fill_f32(float* dst, int64_t n, float value, edl_stream_t stream) {
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_fill_f32: n negative");
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(dst && (((uintptr_t)dst) & 15) == 0, EDL_EINVAL, "edl_fill_f32: dst null or misaligned");
  k_fill<<<grid_for(n / 4 + 1, 1), kThreads, 0, (cudaStream_t)stream>>>(dst, n, value);
  EDL_LAUNCH_CHECK("k_fill");
  return EDL_OK;
}

extern "C" int edl_fill_rows_f32(float* dst, int dim, int64_t row_stride, int64_t n_rows, float value,
                                 edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0 && n_rows >= 0, EDL_EINVAL, "edl_fill_rows_f32: bad dim / n_rows");
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_fill_rows_f32", dim, row_stride, &RS4); if (rc0) return rc0; }
  if (n_rows == 0) return EDL_OK;
  EDL_REQUIRE(dst && (((uintptr_t)dst) & 15) == 0, EDL_EINVAL, "edl_fill_rows_f32: dst null or misaligned");
  k_fill_rows<<<grid_for(n_rows * (dim / 4), 1), kThreads, 0, (cudaStream_t)stream>>>((float4*)dst, dim / 4, RS4, n_rows, value);
  EDL_LAUNCH_CHECK("k_fill_rows");
  return EDL_OK;
}

extern "C" int edl_gather_keys(const float* slab, int dim, int64_t row_stride, const int64_t* keys, int64_t n,
                               const int32_t* n_dev, const int64_t* layout, float* out,
                               edl_stream_t stream) {
  EDL_REQUIRE(n == 0 || keys, EDL_EINVAL, "edl_gather_keys: keys is null");
  KeyRowFn fn{keys};
  return launch_gather<KeyRowFn, true>(slab, dim, row_stride, fn, n, n_dev, layout, out, (cudaStream_t)stream,
                                       "edl_gather_keys");
}

extern "C" int edl_gather_ids(const float* slab, int dim, int64_t row_stride, const int64_t* ids, int64_t n,
                              const int64_t* field_base, int n_fields, float* out,
                              edl_stream_t stream) {
  EDL_REQUIRE(n == 0 || ids, EDL_EINVAL, "edl_gather_ids: ids is null");
  EDL_REQUIRE(!field_base || n_fields > 0, EDL_EINVAL, "edl_gather_ids: n_fields must be > 0 with field_base");
  if (field_base && n_fields >= 2 && n > 0) {
    const uint64_t d = (uint64_t)n_fields;
    const uint64_t M = (1ull << 32) / d + 1;          // < 2^32 for d >= 2
    const uint64_t e = M * d - (1ull << 32);          // 1 <= e <= d
    if ((uint64_t)(n - 1) <= 0xffffffffull / e) {
      IdsRowFnFast fn{ids, field_base, (unsigned)M, (unsigned)n_fields};
      return launch_gather<IdsRowFnFast, false>(slab, dim, row_stride, fn, n, nullptr, nullptr, out, (cudaStream_t)stream,
                                                "edl_gather_ids");
    }
  }
  IdsRowFn fn{ids, field_base, n_fields};
  return launch_gather<IdsRowFn, false>(slab, dim, row_stride, fn, n, nullptr, nullptr, out, (cudaStream_t)stream,
                                        "edl_gather_ids");
}

extern "C" int edl_expand(const float* rows, int dim, const int32_t* index, int64_t n, float* out,
                          edl_stream_t stream) {
  EDL_REQUIRE(n == 0 || index, EDL_EINVAL, "edl_expand: index is null");
  IndexRowFn fn{index};
  return launch_gather<IndexRowFn, false>(rows, dim, 0, fn, n, nullptr, nullptr, out, (cudaStream_t)stream,
                                          "edl_expand");
}

extern "C" int edl_expand_window(const float* rows, int dim, const int32_t* index, int64_t n,
                                 int fields_in_window, int fields_total, int field0, float* out,
                                 edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_expand_window: dim=%d must be a positive multiple of 4", dim);
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_expand_window: n negative");
  EDL_REQUIRE(fields_in_window > 0 && field0 >= 0 && field0 + fields_in_window <= fields_total, EDL_EINVAL,
              "edl_expand_window: window [%d, %d) does not fit %d fields", field0, field0 + fields_in_window,
              fields_total);
  EDL_REQUIRE(n % fields_in_window == 0, EDL_EINVAL, "edl_expand_window: n=%lld is not a multiple of the window's %d fields",
              (long long)n, fields_in_window);
  if (n == 0) return EDL_OK;
  EDL_REQUIRE((n / fields_in_window) * fields_total * (dim / 4) < (int64_t)0xffffffff, EDL_ELIMIT,
              "edl_expand_window: output larger than 2^32 float4 (n=%lld)", (long long)n);
  EDL_REQUIRE(rows && index && out, EDL_EINVAL, "edl_expand_window: null pointer");
  EDL_REQUIRE((((uintptr_t)rows | (uintptr_t)out) & 15) == 0, EDL_EINVAL, "edl_expand_window: rows must be 16-byte aligned");
  const int D4 = dim / 4;
  const int g = grid_for(n * D4, kUnroll);
  cudaStream_t st = (cudaStream_t)stream;
  switch (D4) {
#define EDL_CASE(W) case W: k_expand_window<W><<<g, kThreads, 0, st>>>((const float4*)rows, D4, index, n, fields_in_window, fields_total, field0, (float4*)out); break;
    EDL_CASE(4) EDL_CASE(8) EDL_CASE(16) EDL_CASE(32) EDL_CASE(64)
#undef EDL_CASE
    default: k_expand_window<0><<<g, kThreads, 0, st>>>((const float4*)rows, D4, index, n, fields_in_window, fields_total, field0, (float4*)out);
  }
  EDL_LAUNCH_CHECK("k_expand_window");
  return EDL_OK;
}

extern "C" int edl_compose_index(const int32_t* index, const int32_t* map, int64_t n, int32_t* out,
                                 edl_stream_t stream) {
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_compose_index: n negative");
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(index && map && out, EDL_EINVAL, "edl_compose_index: null pointer");
  k_compose<<<grid_for(n, 1), kThreads, 0, (cudaStream_t)stream>>>(index, map, n, out);
  EDL_LAUNCH_CHECK("k_compose");
  return EDL_OK;
}

extern "C" int edl_scatter_rows(const float* rows, int dim, const int32_t* pos, int64_t n,
                                const int32_t* n_dev, float* out, int64_t out_row_stride, edl_stream_t stream) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "edl_scatter_rows: dim=%d must be a positive multiple of 4", dim);
  int OS4 = 0;
  { int rc0 = edl_row_stride4("edl_scatter_rows", dim, out_row_stride, &OS4); if (rc0) return rc0; }
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_scatter_rows: n negative");
  if (n == 0) return EDL_OK;
  EDL_REQUIRE(rows && pos && out, EDL_EINVAL, "edl_scatter_rows: null pointer");
  const int D4 = dim / 4;
  const int g = grid_for(n * D4, 1);
  cudaStream_t st = (cudaStream_t)stream;
  switch (D4) {
#define EDL_CASE(W) case W: k_scatter_rows<W><<<g, kThreads, 0, st>>>((const float4*)rows, D4, OS4, pos, n, n_dev, (float4*)out); break;
    EDL_CASE(4) EDL_CASE(8) EDL_CASE(16) EDL_CASE(32) EDL_CASE(64)
#undef EDL_CASE
    default: k_scatter_rows<0><<<g, kThreads, 0, st>>>((const float4*)rows, D4, OS4, pos, n, n_dev, (float4*)out);
  }
  EDL_LAUNCH_CHECK("k_scatter_rows");
  return EDL_OK;
}
