// A4 segmented sparse-gradient reduce and A5 Adagrad / Adam row update (fused).
//
// SPEC: g_uniq[u] = sum of the gradient rows that share id u,
// accumulated in float64 and rounded once to fp32; then the optimizer update on the touched
// row with every fp32 operation rounded separately (explicit __f*_rn, so nvcc cannot contract
// to FMA and the result matches the NumPy oracle op for op).
//
// Work split (Zipf skew: a few ids carry thousands of gradient rows):
//   short segments (count <= kLongThreshold): one lane group (dim/4 lanes, <= 32) per segment,
//       kUnroll gradient rows in flight per group;
//   long segments: appended to a list by the short kernel, then one 512-thread CTA per
//       segment sums strided slices in float64 and combines them in a fixed order, so the
//       result is deterministic whatever order the list was filled in.
// HBM traffic per segment: count gradient rows read once (256 B each at dim 64, two full
// 128 B lines), parameter + slot rows read and written once.
#include <algorithm>

#include "edl_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kLongThreads = 512;
constexpr int kUnroll = 4;
constexpr int kLongThreshold = 96;

struct OptDev {
  float lr, eps, b1, b2, omb1, omb2;
};

// ------------------------------------------------------------------ epilogues
struct EpiWrite {  // A4 alone: g_out[dst(u)] = g
  float4* g_out;
  const int32_t* out_pos;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem*, const int64_t*, int u) const {
    return out_pos ? (long long)out_pos[u] : (long long)u;
  }
  __device__ __forceinline__ void prefetch(long long, int, int, float4*, float4*, float4*) const {}
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4 g, float4, float4, float4) const {
    g_out[row * D4 + col] = g;
  }
  __device__ __forceinline__ void touch(long long) const {}
};

__device__ __forceinline__ float adagrad1(float g, float& p, float& a, const OptDev& o) {
  a = __fadd_rn(a, __fmul_rn(g, g));
  float den = __fadd_rn(__fsqrt_rn(a), o.eps);
  p = __fsub_rn(p, __fdiv_rn(__fmul_rn(o.lr, g), den));
  return p;
}

struct EpiAdagrad {
  float4* slab;
  float4* acc;
  const int64_t* uniq;
  uint32_t* touched;
  OptDev o;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem* L, const int64_t* uq, int u) const {
    return edl_key_to_row(L, uq[u]);
  }
  __device__ __forceinline__ void prefetch(long long row, int D4, int col, float4* p, float4* s0, float4*) const {
    *p = slab[row * D4 + col];
    *s0 = acc[row * D4 + col];
  }
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4 g, float4 p, float4 a, float4) const {
    adagrad1(g.x, p.x, a.x, o);
    adagrad1(g.y, p.y, a.y, o);
    adagrad1(g.z, p.z, a.z, o);
    adagrad1(g.w, p.w, a.w, o);
    slab[row * D4 + col] = p;
    acc[row * D4 + col] = a;
  }
  __device__ __forceinline__ void touch(long long row) const {
    if (touched) atomicOr(&touched[row >> 5], 1u << (row & 31));
  }
};

__device__ __forceinline__ void adam1(float g, float& p, float& m, float& v, const OptDev& o) {
  m = __fadd_rn(__fmul_rn(o.b1, m), __fmul_rn(o.omb1, g));
  v = __fadd_rn(__fmul_rn(o.b2, v), __fmul_rn(__fmul_rn(o.omb2, g), g));
  float den = __fadd_rn(__fsqrt_rn(v), o.eps);
  p = __fsub_rn(p, __fdiv_rn(__fmul_rn(o.lr, m), den));
}

struct EpiAdam {
  float4* slab;
  float4* m;
  float4* v;
  const int64_t* uniq;
  uint32_t* touched;
  OptDev o;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem* L, const int64_t* uq, int u) const {
    return edl_key_to_row(L, uq[u]);
  }
  __device__ __forceinline__ void prefetch(long long row, int D4, int col, float4* p, float4* s0, float4* s1) const {
    *p = slab[row * D4 + col];
    *s0 = m[row * D4 + col];
    *s1 = v[row * D4 + col];
  }
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4 g, float4 p, float4 mm, float4 vv) const {
    adam1(g.x, p.x, mm.x, vv.x, o);
    adam1(g.y, p.y, mm.y, vv.y, o);
    adam1(g.z, p.z, mm.z, vv.z, o);
    adam1(g.w, p.w, mm.w, vv.w, o);
    slab[row * D4 + col] = p;
    m[row * D4 + col] = mm;
    v[row * D4 + col] = vv;
  }
  __device__ __forceinline__ void touch(long long row) const {
    if (touched) atomicOr(&touched[row >> 5], 1u << (row & 31));
  }
};

// ------------------------------------------------------------------ short segments
// kG lanes per segment, kC float4 chunks per lane: lane gl owns columns gl + c*kG < D4.
template <int kG, int kC, typename Epi, bool kUseLayout>
__global__ void __launch_bounds__(kThreads)
k_seg_short(const float4* __restrict__ grad, int D4, const int32_t* __restrict__ order,
            const int32_t* __restrict__ seg, const int32_t* __restrict__ n_unique,
            const int64_t* __restrict__ uniq, const int64_t* __restrict__ layout, Epi epi,
            int32_t* __restrict__ long_count, int32_t* __restrict__ long_list) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, kUseLayout ? layout : nullptr);
  __syncthreads();
  const int U = *n_unique;
  const int u = (int)(((int64_t)blockIdx.x * kThreads + threadIdx.x) / kG);
  const int gl = threadIdx.x % kG;
  if (u >= U) return;
  const int s0 = seg[u], s1 = seg[u + 1];
  const int cnt = s1 - s0;
  if (cnt > kLongThreshold) {
    if (gl == 0) long_list[atomicAdd(long_count, 1)] = u;
    return;
  }
  const long long row = epi.row_of(&L, uniq, u);
  float4 p[kC], a0[kC], a1[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) {
    int col = gl + c * kG;
    if (col < D4) epi.prefetch(row, D4, col, &p[c], &a0[c], &a1[c]);
  }
  float4 g[kC];
  if (cnt == 1) {
    const long long j = order[s0];
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      int col = gl + c * kG;
      if (col < D4) g[c] = grad[j * D4 + col];
    }
  } else {
    double acc[kC][4];
#pragma unroll
    for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
    for (int i = s0; i < s1; i += kUnroll) {
      float4 v[kUnroll][kC];
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        if (i + k < s1) {
          const long long j = order[i + k];
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            int col = gl + c * kG;
            if (col < D4) v[k][c] = grad[j * D4 + col];
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        if (i + k < s1) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            acc[c][0] += (double)v[k][c].x;
            acc[c][1] += (double)v[k][c].y;
            acc[c][2] += (double)v[k][c].z;
            acc[c][3] += (double)v[k][c].w;
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kC; ++c)
      g[c] = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                         __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
  }
#pragma unroll
  for (int c = 0; c < kC; ++c) {
    int col = gl + c * kG;
    if (col < D4) epi.apply(row, D4, col, g[c], p[c], a0[c], a1[c]);
  }
  if (gl == 0) epi.touch(row);
}

// ------------------------------------------------------------------ long segments
// One CTA per long segment.  Lane group k sums rows s0 + k, s0 + k + NG, ... in float64;
// the NG partial rows are then added in index order by the threads that own the columns.
template <int kG, int kC, typename Epi, bool kUseLayout>
__global__ void __launch_bounds__(kLongThreads)
k_seg_long(const float4* __restrict__ grad, int D4, const int32_t* __restrict__ order,
           const int32_t* __restrict__ seg, const int64_t* __restrict__ uniq,
           const int64_t* __restrict__ layout, Epi epi, const int32_t* __restrict__ long_count,
           const int32_t* __restrict__ long_list) {
  extern __shared__ double part[];  // [NG][D4*4]
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, kUseLayout ? layout : nullptr);
  __syncthreads();
  constexpr int NG = kLongThreads / kG;
  const int gk = threadIdx.x / kG, gl = threadIdx.x % kG;
  const int D = D4 * 4;
  const int n_long = *long_count;
  for (int li = blockIdx.x; li < n_long; li += gridDim.x) {
    const int u = long_list[li];
    const int s0 = seg[u], s1 = seg[u + 1];
    double acc[kC][4];
#pragma unroll
    for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
    for (int i = s0 + gk; i < s1; i += NG * kUnroll) {
      float4 v[kUnroll][kC];
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        int ii = i + k * NG;
        if (ii < s1) {
          const long long j = order[ii];
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            int col = gl + c * kG;
            if (col < D4) v[k][c] = grad[j * D4 + col];
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        int ii = i + k * NG;
        if (ii < s1) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            acc[c][0] += (double)v[k][c].x;
            acc[c][1] += (double)v[k][c].y;
            acc[c][2] += (double)v[k][c].z;
            acc[c][3] += (double)v[k][c].w;
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      int col = gl + c * kG;
      if (col < D4) {
        double* dst = part + (size_t)gk * D + col * 4;
        dst[0] = acc[c][0]; dst[1] = acc[c][1]; dst[2] = acc[c][2]; dst[3] = acc[c][3];
      }
    }
    __syncthreads();
    if ((int)threadIdx.x < D4) {
      const int col = threadIdx.x;
      double s[4] = {0.0, 0.0, 0.0, 0.0};
      for (int k = 0; k < NG; ++k) {
        const double* src = part + (size_t)k * D + col * 4;
        s[0] += src[0]; s[1] += src[1]; s[2] += src[2]; s[3] += src[3];
      }
      const long long row = epi.row_of(&L, uniq, u);
      float4 p, a0, a1;
      epi.prefetch(row, D4, col, &p, &a0, &a1);
      float4 g = make_float4(__double2float_rn(s[0]), __double2float_rn(s[1]),
                             __double2float_rn(s[2]), __double2float_rn(s[3]));
      epi.apply(row, D4, col, g, p, a0, a1);
      if (col == 0) epi.touch(row);
    }
    __syncthreads();
  }
}

struct SegWs { int32_t* long_count; int32_t* long_list; size_t bytes; };
SegWs carve_seg(void* ws, int64_t max_unique) {
  SegWs r; char* p = (char*)ws;
  r.long_count = (int32_t*)p;
  r.long_list = (int32_t*)(p + 256);
  r.bytes = 256 + edl_align_up((size_t)(max_unique > 0 ? max_unique : 1) * 4, 256);
  return r;
}

template <int kG, int kC, typename Epi, bool kUseLayout>
int launch_cfg(const float* grad, int D4, const int32_t* order, const int32_t* seg,
               const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
               const int64_t* layout, Epi epi, const SegWs& W, cudaStream_t st) {
  EDL_CUDA(cudaMemsetAsync(W.long_count, 0, sizeof(int32_t), st));
  const int64_t ctas = edl_cdiv(max_unique * kG, kThreads);
  EDL_REQUIRE(ctas < 0x7fffffff, EDL_ELIMIT, "segreduce: max_unique=%lld too large", (long long)max_unique);
  k_seg_short<kG, kC, Epi, kUseLayout><<<(unsigned)ctas, kThreads, 0, st>>>(
      (const float4*)grad, D4, order, seg, n_unique, uniq, layout, epi, W.long_count, W.long_list);
  EDL_LAUNCH_CHECK("k_seg_short");
  const size_t smem = (size_t)(kLongThreads / kG) * D4 * 4 * sizeof(double);
  if (smem > 48 * 1024) {
    EDL_CUDA(cudaFuncSetAttribute(k_seg_long<kG, kC, Epi, kUseLayout>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  const int g = (int)std::min<int64_t>(std::max<int64_t>(max_unique / kLongThreshold, 1), EDL_SM_COUNT * 2);
  k_seg_long<kG, kC, Epi, kUseLayout><<<g, kLongThreads, smem, st>>>(
      (const float4*)grad, D4, order, seg, uniq, layout, epi, W.long_count, W.long_list);
  EDL_LAUNCH_CHECK("k_seg_long");
  return EDL_OK;
}

template <typename Epi, bool kUseLayout>
int dispatch_dim(const float* grad, int dim, const int32_t* order, const int32_t* seg,
                 const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
                 const int64_t* layout, Epi epi, const SegWs& W, cudaStream_t st) {
  const int D4 = dim / 4;
#define EDL_GO(G, C) return launch_cfg<G, C, Epi, kUseLayout>(grad, D4, order, seg, n_unique, max_unique, uniq, layout, epi, W, st)
  if (D4 <= 4) EDL_GO(4, 1);
  if (D4 <= 8) EDL_GO(8, 1);
  if (D4 <= 16) EDL_GO(16, 1);
  if (D4 <= 32) EDL_GO(32, 1);
  if (D4 <= 64) EDL_GO(32, 2);
  if (D4 <= 128) EDL_GO(32, 4);
#undef EDL_GO
  edl_set_error("segreduce: dim=%d above the supported maximum 512", dim);
  return EDL_ELIMIT;
}

int check_common(const char* who, const float* grad, int dim, const int32_t* order,
                 const int32_t* seg, const int32_t* n_unique, int64_t max_unique, void* ws,
                 size_t ws_bytes, SegWs* W) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "%s: dim=%d must be a positive multiple of 4", who, dim);
  EDL_REQUIRE(max_unique >= 0, EDL_EINVAL, "%s: max_unique negative", who);
  EDL_REQUIRE(grad && order && seg && n_unique, EDL_EINVAL, "%s: null pointer", who);
  EDL_REQUIRE((((uintptr_t)grad) & 15) == 0, EDL_EINVAL, "%s: grad must be 16-byte aligned", who);
  *W = carve_seg(ws, max_unique);
  EDL_REQUIRE(ws && ws_bytes >= W->bytes, EDL_EWORKSPACE, "%s: workspace %zu < required %zu", who,
              ws_bytes, W->bytes);
  return EDL_OK;
}

}  // namespace

extern "C" size_t edl_segreduce_workspace_bytes(int64_t max_unique, int dim) {
  (void)dim;
  return carve_seg(nullptr, max_unique).bytes;
}

extern "C" int edl_segreduce(const float* grad, int dim, const int32_t* order,
                             const int32_t* seg_offsets, const int32_t* n_unique_dev,
                             int64_t max_unique, const int32_t* out_pos, float* g_out,
                             void* workspace, size_t workspace_bytes, edl_stream_t stream) {
  if (max_unique == 0) return EDL_OK;
  SegWs W;
  int rc = check_common("edl_segreduce", grad, dim, order, seg_offsets, n_unique_dev, max_unique,
                        workspace, workspace_bytes, &W);
  if (rc) return rc;
  EDL_REQUIRE(g_out && (((uintptr_t)g_out) & 15) == 0, EDL_EINVAL, "edl_segreduce: g_out null or misaligned");
  EpiWrite epi{(float4*)g_out, out_pos};
  return dispatch_dim<EpiWrite, false>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, nullptr,
                                       nullptr, epi, W, (cudaStream_t)stream);
}

extern "C" int edl_segreduce_update(float* slab, float* slot0, float* slot1, int dim,
                                    const int64_t* uniq, const int32_t* order,
                                    const int32_t* seg_offsets, const int32_t* n_unique_dev,
                                    int64_t max_unique, const float* grad, const int64_t* layout,
                                    const edl_opt_params* opt, uint32_t* touched, void* workspace,
                                    size_t workspace_bytes, edl_stream_t stream) {
  if (max_unique == 0) return EDL_OK;
  SegWs W;
  int rc = check_common("edl_segreduce_update", grad, dim, order, seg_offsets, n_unique_dev,
                        max_unique, workspace, workspace_bytes, &W);
  if (rc) return rc;
  EDL_REQUIRE(opt, EDL_EINVAL, "edl_segreduce_update: opt is null");
  EDL_REQUIRE(slab && slot0 && uniq, EDL_EINVAL, "edl_segreduce_update: null slab / slot0 / uniq");
  EDL_REQUIRE((((uintptr_t)slab | (uintptr_t)slot0 | (uintptr_t)slot1) & 15) == 0, EDL_EINVAL,
              "edl_segreduce_update: rows must be 16-byte aligned");
  OptDev o;
  o.lr = opt->lr; o.eps = opt->eps; o.b1 = opt->beta1; o.b2 = opt->beta2;
  o.omb1 = 1.0f - opt->beta1;  // fp32 subtraction, as the oracle does
  o.omb2 = 1.0f - opt->beta2;
  cudaStream_t st = (cudaStream_t)stream;
  if (opt->kind == EDL_OPT_ADAGRAD) {
    EpiAdagrad epi{(float4*)slab, (float4*)slot0, uniq, touched, o};
    return layout ? dispatch_dim<EpiAdagrad, true>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st)
                  : dispatch_dim<EpiAdagrad, false>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st);
  }
  if (opt->kind == EDL_OPT_ADAM) {
    EDL_REQUIRE(slot1, EDL_EINVAL, "edl_segreduce_update: Adam needs slot1 (v)");
    EpiAdam epi{(float4*)slab, (float4*)slot0, (float4*)slot1, uniq, touched, o};
    return layout ? dispatch_dim<EpiAdam, true>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st)
                  : dispatch_dim<EpiAdam, false>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st);
  }
  edl_set_error("edl_segreduce_update: unknown optimizer kind %d", opt->kind);
  return EDL_EINVAL;
}
