// rowbw_probe: what HBM delivers for the ACCESS PATTERNS of the sparse step, with the arithmetic and the
// index work taken out -- the ceiling the lookup / update kernels can be held against, next to the
// sequential-copy peak of MEASURED_PEAKS.json.
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/rowbw_probe tools/rowbw_probe.cu
//   tools/rowbw_probe [table_rows=260000000] [dim=64] [n_lookup=1703936] [n_unique=708000] [reps=20]
//
// Patterns (rows of dim fp32; row indices are precomputed int32, uniformly random, all distinct):
//   copy        out[i] = in[i]                                   sequential read + write (calibration)
//   gather      out[i] = slab[idx[i]]           n_lookup rows    random row reads, sequential writes (the lookup)
//   gather_u    the same with n_unique distinct rows repeated to n_lookup (Zipf-like re-use through L2)
//   update      slab[r] += g[i]; acc[r] += g[i]  n_unique rows   3 random row reads + 2 random row writes (the update)
//   update_ef   the same with evict-first loads / stores for g and acc, as the product kernels use
//   update_interleaved  the same with parameter and accumulator side by side in 512-byte records
//   read3 / write2      the update's reads alone / its writes alone
// One thread per float4, 8 CTAs x 256 threads per SM, two rows in flight per thread.  Prints GB/s per pattern
// (bytes = what the pattern must move) as one JSON object.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e = (x);                                                               \
    if (e != cudaSuccess) {                                                            \
      fprintf(stderr, "%s: %s (%s:%d)\n", #x, cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                         \
    }                                                                                  \
  } while (0)

constexpr int kT = 256;

__global__ void __launch_bounds__(kT, 8) k_copy(const float4* __restrict__ in, float4* __restrict__ out, int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * kT;
  for (int64_t e = (int64_t)blockIdx.x * kT + threadIdx.x; e < n4; e += 2 * stride) {
    float4 a = in[e], b = make_float4(0, 0, 0, 0);
    if (e + stride < n4) b = in[e + stride];
    out[e] = a;
    if (e + stride < n4) out[e + stride] = b;
  }
}

template <int D4>
__global__ void __launch_bounds__(kT, 8)
k_gather(const float4* __restrict__ slab, const int32_t* __restrict__ idx, int64_t n, float4* __restrict__ out) {
  const int64_t total = n * D4, stride = (int64_t)gridDim.x * kT;
  const int c = threadIdx.x % D4;
  for (int64_t e = (int64_t)blockIdx.x * kT + threadIdx.x; e < total; e += 2 * stride) {
    const int64_t e1 = e + stride;
    const int64_t r0 = idx[e / D4], r1 = e1 < total ? idx[e1 / D4] : 0;
    float4 a = slab[r0 * D4 + c], b = make_float4(0, 0, 0, 0);
    if (e1 < total) b = slab[r1 * D4 + c];
    __stcs(out + e, a);
    if (e1 < total) __stcs(out + e1, b);
  }
}

template <int D4, bool kEF>
__global__ void __launch_bounds__(kT, 8)
k_update(float4* __restrict__ slab, float4* __restrict__ acc, const float4* __restrict__ g,
         const int32_t* __restrict__ idx, const int32_t* __restrict__ gidx, int64_t n) {
  const int64_t total = n * D4, stride = (int64_t)gridDim.x * kT;
  const int c = threadIdx.x % D4;
  for (int64_t e = (int64_t)blockIdx.x * kT + threadIdx.x; e < total; e += stride) {
    const int64_t i = e / D4;
    const int64_t r = idx[i], j = gidx[i];
    float4 p = kEF ? __ldcs(slab + r * D4 + c) : slab[r * D4 + c];
    float4 a = kEF ? __ldcs(acc + r * D4 + c) : acc[r * D4 + c];
    float4 x = kEF ? __ldcs(g + j * D4 + c) : g[j * D4 + c];
    p.x += x.x; p.y += x.y; p.z += x.z; p.w += x.w;
    a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    slab[r * D4 + c] = p;
    if (kEF) __stcs(acc + r * D4 + c, a); else acc[r * D4 + c] = a;
  }
}

// the same update with parameter and accumulator of a row side by side (512-byte records [p | a])
template <int D4>
__global__ void __launch_bounds__(kT, 8)
k_update_aos(float4* __restrict__ rec, const float4* __restrict__ g, const int32_t* __restrict__ idx,
             const int32_t* __restrict__ gidx, int64_t n) {
  const int64_t total = n * D4, stride = (int64_t)gridDim.x * kT;
  const int c = threadIdx.x % D4;
  for (int64_t e = (int64_t)blockIdx.x * kT + threadIdx.x; e < total; e += stride) {
    const int64_t i = e / D4;
    const int64_t r = idx[i], j = gidx[i];
    float4 p = rec[r * 2 * D4 + c], a = rec[r * 2 * D4 + D4 + c], x = g[j * D4 + c];
    p.x += x.x; p.y += x.y; p.z += x.z; p.w += x.w;
    a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    rec[r * 2 * D4 + c] = p;
    rec[r * 2 * D4 + D4 + c] = a;
  }
}
// reads only (3 random rows per id, one float4 of the sum written per id) / writes only (2 random rows per id)
template <int D4>
__global__ void __launch_bounds__(kT, 8)
k_read3(const float4* __restrict__ slab, const float4* __restrict__ acc, const float4* __restrict__ g,
        const int32_t* __restrict__ idx, const int32_t* __restrict__ gidx, int64_t n, float4* __restrict__ out) {
  const int64_t total = n * D4, stride = (int64_t)gridDim.x * kT;
  const int c = threadIdx.x % D4;
  for (int64_t e = (int64_t)blockIdx.x * kT + threadIdx.x; e < total; e += stride) {
    const int64_t i = e / D4;
    const int64_t r = idx[i], j = gidx[i];
    float4 p = slab[r * D4 + c], a = acc[r * D4 + c], x = g[j * D4 + c];
    p.x += x.x + a.x; p.y += x.y + a.y; p.z += x.z + a.z; p.w += x.w + a.w;
    if (p.x == 12345.678f) out[i] = p;   // never true: keeps the loads alive
  }
}
template <int D4>
__global__ void __launch_bounds__(kT, 8)
k_write2(float4* __restrict__ slab, float4* __restrict__ acc, const int32_t* __restrict__ idx, int64_t n) {
  const int64_t total = n * D4, stride = (int64_t)gridDim.x * kT;
  const int c = threadIdx.x % D4;
  for (int64_t e = (int64_t)blockIdx.x * kT + threadIdx.x; e < total; e += stride) {
    const int64_t r = idx[e / D4];
    const float4 v = make_float4(1.f, 2.f, 3.f, (float)c);
    slab[r * D4 + c] = v;
    acc[r * D4 + c] = v;
  }
}

__global__ void k_fill_idx(int32_t* idx, int64_t n, int64_t rows, uint64_t seed) {
  // distinct pseudo-random rows: a stride permutation of [0, rows) (rows coprime with the multiplier)
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    idx[i] = (int32_t)(((unsigned __int128)(i + seed) * 2654435761ull) % (uint64_t)rows);
}
__global__ void k_repeat_idx(int32_t* out, const int32_t* uniq, int64_t n, int64_t u) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = uniq[(i * 2654435761ull) % (uint64_t)u];
}

template <typename F>
double time_ms(F f, int reps) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) f(i);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < reps; ++i) f(3 + i);
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  CK(cudaGetLastError());
  return ms / reps;
}

int main(int argc, char** argv) {
  const int64_t rows = argc > 1 ? atoll(argv[1]) : 260000000ll;
  const int dim = argc > 2 ? atoi(argv[2]) : 64;
  const int64_t n_lookup = argc > 3 ? atoll(argv[3]) : 1703936;
  const int64_t n_unique = argc > 4 ? atoll(argv[4]) : 708000;
  const int reps = argc > 5 ? atoi(argv[5]) : 20;
  if (dim != 64) { fprintf(stderr, "this probe is built for dim 64\n"); return 2; }
  constexpr int D4 = 16;
  const int grid = 148 * 8;
  const int NB = 4;   // distinct index sets cycled through, so no repetition is served from L2
  float4 *slab, *acc, *g, *out;
  int32_t *idx, *idx_u, *idx_rep, *gidx;
  CK(cudaMalloc(&slab, (size_t)rows * dim * 4));
  CK(cudaMalloc(&acc, (size_t)rows * dim * 4));
  CK(cudaMalloc(&g, (size_t)n_lookup * dim * 4));
  CK(cudaMalloc(&out, (size_t)n_lookup * dim * 4));
  CK(cudaMalloc(&idx, (size_t)NB * n_lookup * 4));
  CK(cudaMalloc(&idx_u, (size_t)NB * n_unique * 4));
  CK(cudaMalloc(&idx_rep, (size_t)NB * n_lookup * 4));
  CK(cudaMalloc(&gidx, (size_t)n_lookup * 4));
  CK(cudaMemset(slab, 0, (size_t)rows * dim * 4));
  CK(cudaMemset(acc, 0, (size_t)rows * dim * 4));
  CK(cudaMemset(g, 0, (size_t)n_lookup * dim * 4));
  for (int b = 0; b < NB; ++b) {
    k_fill_idx<<<1184, 256>>>(idx + b * n_lookup, n_lookup, rows, 1000003ull * (b + 1));
    k_fill_idx<<<1184, 256>>>(idx_u + b * n_unique, n_unique, rows, 7000003ull * (b + 1));
    k_repeat_idx<<<1184, 256>>>(idx_rep + b * n_lookup, idx_u + b * n_unique, n_lookup, n_unique);
  }
  k_fill_idx<<<1184, 256>>>(gidx, n_lookup, n_lookup, 12345);
  CK(cudaDeviceSynchronize());

  const int64_t copy4 = (int64_t)1 << 28;   // 4 GiB in, 4 GiB out (inside the slab / acc allocations)
  const double t_copy = time_ms([&](int) { k_copy<<<grid, kT>>>(slab, acc, copy4); }, reps);
  const double t_gather = time_ms([&](int i) { k_gather<D4><<<grid, kT>>>(slab, idx + (i % NB) * n_lookup, n_lookup, out); }, reps);
  const double t_gather_u = time_ms([&](int i) { k_gather<D4><<<grid, kT>>>(slab, idx_rep + (i % NB) * n_lookup, n_lookup, out); }, reps);
  const double t_upd = time_ms([&](int i) { k_update<D4, false><<<grid, kT>>>(slab, acc, g, idx_u + (i % NB) * n_unique, gidx, n_unique); }, reps);
  const double t_upd_ef = time_ms([&](int i) { k_update<D4, true><<<grid, kT>>>(slab, acc, g, idx_u + (i % NB) * n_unique, gidx, n_unique); }, reps);
  // interleaved records live in the first half of the slab allocation (rows / 2 records of 512 bytes)
  int32_t* idx_h;
  CK(cudaMalloc(&idx_h, (size_t)NB * n_unique * 4));
  for (int b = 0; b < NB; ++b) k_fill_idx<<<1184, 256>>>(idx_h + b * n_unique, n_unique, rows / 2, 9000011ull * (b + 1));
  CK(cudaDeviceSynchronize());
  const double t_aos = time_ms([&](int i) { k_update_aos<D4><<<grid, kT>>>(slab, g, idx_h + (i % NB) * n_unique, gidx, n_unique); }, reps);
  const double t_r3 = time_ms([&](int i) { k_read3<D4><<<grid, kT>>>(slab, acc, g, idx_u + (i % NB) * n_unique, gidx, n_unique, out); }, reps);
  const double t_w2 = time_ms([&](int i) { k_write2<D4><<<grid, kT>>>(slab, acc, idx_u + (i % NB) * n_unique, n_unique); }, reps);
  const double rowb = dim * 4.0;
  const double b_copy = 2.0 * copy4 * 16, b_gather = n_lookup * (2 * rowb + 4), b_gather_u = n_lookup * (rowb + 4) + n_unique * rowb,
               b_upd = n_unique * (5 * rowb + 8);
  printf("{\"table_rows\": %lld, \"dim\": %d, \"n_lookup\": %lld, \"n_unique\": %lld, "
         "\"copy_gbs\": %.1f, \"gather_distinct_gbs\": %.1f, \"gather_reuse_gbs\": %.1f, \"update_gbs\": %.1f, \"update_evict_first_gbs\": %.1f, "
         "\"update_interleaved_gbs\": %.1f, \"read3_gbs\": %.1f, \"write2_gbs\": %.1f, "
         "\"update_interleaved_ms\": %.4f, \"read3_ms\": %.4f, \"write2_ms\": %.4f, "
         "\"copy_ms\": %.4f, \"gather_distinct_ms\": %.4f, \"gather_reuse_ms\": %.4f, \"update_ms\": %.4f, \"update_evict_first_ms\": %.4f}\n",
         (long long)rows, dim, (long long)n_lookup, (long long)n_unique, b_copy / t_copy / 1e6, b_gather / t_gather / 1e6,
         b_gather_u / t_gather_u / 1e6, b_upd / t_upd / 1e6, b_upd / t_upd_ef / 1e6,
         b_upd / t_aos / 1e6, n_unique * (3 * rowb + 8) / t_r3 / 1e6, n_unique * (2 * rowb + 4) / t_w2 / 1e6, t_aos, t_r3, t_w2, t_copy, t_gather, t_gather_u, t_upd, t_upd_ef);
  return 0;
}
