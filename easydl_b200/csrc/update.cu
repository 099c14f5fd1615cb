// A4 segmented sparse-gradient reduce and A5 Adagrad / Adam row update (fused).
//
// SPEC: g_uniq[u] = sum of the gradient rows that share id u,
// accumulated in float64 and rounded once to fp32; then the optimizer update on the touched
// row with every fp32 operation rounded separately (explicit __f*_rn, so nvcc cannot contract
// to FMA and the result matches the NumPy oracle op for op).
//
// Kernel design (k_seg_tile): HBM-bound gather-reduce-scatter, so the only thing that matters
// is bytes in flight.  A persistent CTA per SM slot owns a contiguous slice of the SORTED
// positions (perfect balance whatever the Zipf skew: a hot id with 100K gradient rows is just
// more positions).  It walks the slice in stages of <= ROWS positions / <= SEGS segments:
//   producer side: every thread issues one TMA bulk copy (cp.async.bulk global -> shared,
//       completion on an mbarrier) for "its" gradient row grad[order[p + tid]], and, for each
//       segment that finishes inside the stage, for its parameter row and slot row(s);
//   consumer side: lane groups (dim/4 lanes) sum a segment's rows from shared memory in
//       float64 and run the optimizer epilogue straight out of shared memory.
// Two stage buffers: the copies of stage k+1 are in flight while stage k is reduced, ~100 KB
// of HBM reads outstanding per SM (latency x bandwidth needs ~45 KB).  Segments that cross a
// stage boundary carry a float64 partial row in shared memory; segments that cross a CTA
// boundary leave float64 partials in global scratch and the last CTA to finish (atomic
// ticket, no spinning) adds them in CTA order and applies the update -- deterministic.
#include <algorithm>

#include "edl_common.cuh"

namespace {

constexpr int kThreads = 256;

struct OptDev {
  float lr, eps, b1, b2, omb1, omb2;
};

// ------------------------------------------------------------------ PTX helpers (sm_90+)
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// TMA bulk copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, unsigned bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ------------------------------------------------------------------ gradient row element types
template <typename GT> struct Row;
template <> struct Row<float> {
  typedef float4 V;
  static constexpr int kBytes4 = 16;  // bytes of 4 consecutive elements
  static __device__ __forceinline__ V ld(const unsigned char* row, int col) { return *(const float4*)(row + col * 16); }
  static __device__ __forceinline__ void add(double* a, V v) { a[0] += (double)v.x; a[1] += (double)v.y; a[2] += (double)v.z; a[3] += (double)v.w; }
  static __device__ __forceinline__ float4 f32(V v) { return v; }
};
template <> struct Row<double> {
  struct V { double2 a, b; };
  static constexpr int kBytes4 = 32;
  static __device__ __forceinline__ V ld(const unsigned char* row, int col) {
    V v; v.a = *(const double2*)(row + col * 32); v.b = *(const double2*)(row + col * 32 + 16); return v;
  }
  static __device__ __forceinline__ void add(double* a, V v) { a[0] += v.a.x; a[1] += v.a.y; a[2] += v.b.x; a[3] += v.b.y; }
  static __device__ __forceinline__ float4 f32(V v) {
    return make_float4(__double2float_rn(v.a.x), __double2float_rn(v.a.y), __double2float_rn(v.b.x), __double2float_rn(v.b.y));
  }
};

// ------------------------------------------------------------------ epilogues
// kSlots = number of fp32 row arrays staged per finishing segment (parameter row + slots).
struct EpiWrite {  // A4 alone: g_out[dst(u)] = g (fp32 rows)
  static constexpr int kSlots = 0;
  static constexpr bool kOutF64 = false;
  float4* g_out;
  const int32_t* out_pos;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem*, const int64_t*, int u) const {
    return out_pos ? (long long)out_pos[u] : (long long)u;
  }
  __device__ __forceinline__ const float* base(int) const { return nullptr; }
  __device__ __forceinline__ void load(long long, int, int, float4*, float4*, float4*) const {}
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4 g, const double*, float4, float4, float4) const {
    g_out[row * D4 + col] = g;
  }
  __device__ __forceinline__ void touch(long long) const {}
};

struct EpiWrite64 {  // A4 alone, float64 rows out: the exact wire format of the gradient exchange
  static constexpr int kSlots = 0;
  static constexpr bool kOutF64 = true;
  double* g_out;
  const int32_t* out_pos;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem*, const int64_t*, int u) const {
    return out_pos ? (long long)out_pos[u] : (long long)u;
  }
  __device__ __forceinline__ const float* base(int) const { return nullptr; }
  __device__ __forceinline__ void load(long long, int, int, float4*, float4*, float4*) const {}
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4, const double* s, float4, float4, float4) const {
    double2* dst = (double2*)(g_out + (row * D4 + col) * 4);
    dst[0] = make_double2(s[0], s[1]);
    dst[1] = make_double2(s[2], s[3]);
  }
  __device__ __forceinline__ void touch(long long) const {}
};

__device__ __forceinline__ void adagrad1(float g, float& p, float& a, const OptDev& o) {
  a = __fadd_rn(a, __fmul_rn(g, g));
  float den = __fadd_rn(__fsqrt_rn(a), o.eps);
  p = __fsub_rn(p, __fdiv_rn(__fmul_rn(o.lr, g), den));
}

struct EpiAdagrad {
  static constexpr int kSlots = 2;
  static constexpr bool kOutF64 = false;
  float4* slab;
  float4* acc;
  uint32_t* touched;
  OptDev o;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem* L, const int64_t* uq, int u) const {
    return edl_key_to_row(L, uq[u]);
  }
  __device__ __forceinline__ const float* base(int s) const { return s == 0 ? (const float*)slab : (const float*)acc; }
  __device__ __forceinline__ void load(long long row, int D4, int col, float4* p, float4* s0, float4*) const {
    *p = slab[row * D4 + col];
    *s0 = acc[row * D4 + col];
  }
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4 g, const double*, float4 p, float4 a, float4) const {
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
  static constexpr int kSlots = 3;
  static constexpr bool kOutF64 = false;
  float4* slab;
  float4* m;
  float4* v;
  uint32_t* touched;
  OptDev o;
  __device__ __forceinline__ long long row_of(const EdlLayoutSmem* L, const int64_t* uq, int u) const {
    return edl_key_to_row(L, uq[u]);
  }
  __device__ __forceinline__ const float* base(int s) const {
    return s == 0 ? (const float*)slab : (s == 1 ? (const float*)m : (const float*)v);
  }
  __device__ __forceinline__ void load(long long row, int D4, int col, float4* p, float4* s0, float4* s1) const {
    *p = slab[row * D4 + col];
    *s0 = m[row * D4 + col];
    *s1 = v[row * D4 + col];
  }
  __device__ __forceinline__ void apply(long long row, int D4, int col, float4 g, const double*, float4 p, float4 mm, float4 vv) const {
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

// ------------------------------------------------------------------ the tile kernel
struct TileParams {
  const unsigned char* grad;  // [N, D] rows of GT
  int D4;
  const int32_t* order;
  const int32_t* seg;
  const int32_t* n_unique;
  const int64_t* uniq;
  const int64_t* layout;
  int rows_per_stage;  // <= kThreads
  int segs_per_stage;  // <= kThreads
  double* carry_h;     // [grid][D]  partial of the segment that continues from the previous CTA
  double* carry_t;     // [grid][D]  partial of the last segment when it continues into the next CTA
  int32_t* flags;      // [grid][4]  h_valid, h_ends, t_valid, t_seg
  unsigned* ticket;
};

template <int kG, int kC, typename GT, typename Epi>
__global__ void __launch_bounds__(kThreads)
k_seg_tile(TileParams P, Epi epi) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ EdlLayoutSmem L;
  __shared__ uint64_t bar[2];
  __shared__ int s_last;
  constexpr int NG = kThreads / kG;
  constexpr int NS = Epi::kSlots;
  typedef Row<GT> R;
  const int tid = threadIdx.x;
  const int gk = tid / kG, gl = tid % kG;
  const int D4 = P.D4, D = D4 * 4;
  const int ROWS = P.rows_per_stage, SEGS = P.segs_per_stage;
  const unsigned ROWB = (unsigned)D4 * 16u;              // fp32 row bytes
  const unsigned ROWB_G = (unsigned)D4 * R::kBytes4;     // gradient row bytes
  const size_t stage_bytes = (size_t)ROWS * ROWB_G + (size_t)NS * SEGS * ROWB;
  double* carry_s = (double*)(smem + 2 * stage_bytes);   // [2][D]
  int* segs_s = (int*)(carry_s + 2 * D);                 // [2][SEGS + 2]

  edl_layout_load(&L, P.layout);
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_proxy_async();
  }
  __syncthreads();

  const int U = *P.n_unique;
  const int N = U > 0 ? P.seg[U] : 0;
  const int per = (int)(((long long)N + gridDim.x - 1) / gridDim.x);
  const int P0 = min(N, (int)blockIdx.x * per);
  const int P1 = min(N, P0 + per);

  if (P0 < P1) {
    // segment containing P0: largest u with seg[u] <= P0
    int lo = 0, hi = U;
    while (hi - lo > 1) {
      int mid = (lo + hi) >> 1;
      if (P.seg[mid] <= P0) lo = mid; else hi = mid;
    }
    const int u0 = lo;
    const bool head = P.seg[u0] < P0;  // u0 started in an earlier CTA's slice

    // ---- stage builder: loads segment offsets, decides the stage extent, issues the copies
    auto build = [&](int b, int p, int u, int& q_out, int& nseg_out) {
      int* sg = segs_s + b * (SEGS + 2);
      for (int i = tid; i <= SEGS; i += kThreads) sg[i] = P.seg[min(u + i, U)];
      __syncthreads();
      const int q = min(min(p + ROWS, P1), sg[SEGS]);
      const int nseg = __syncthreads_count(tid < SEGS && sg[tid] < q);
      bool fin = false;
      if (NS > 0) fin = tid < nseg && sg[tid + 1] <= q && !(head && u + tid == u0);
      const int nfin = NS > 0 ? __syncthreads_count(fin) : 0;
      if (tid == 0) mbar_expect_tx(&bar[b], (unsigned)(q - p) * ROWB_G + (unsigned)nfin * ROWB * NS);
      if (tid < q - p) {
        const long long j = P.order[p + tid];
        bulk_g2s((smem + (size_t)b * stage_bytes) + (size_t)tid * ROWB_G, P.grad + (size_t)j * ROWB_G, ROWB_G, &bar[b]);
      }
      if (NS > 0 && fin) {
        const long long row = epi.row_of(&L, P.uniq, u + tid);
        unsigned char* dst = (smem + (size_t)b * stage_bytes) + (size_t)ROWS * ROWB_G + (size_t)tid * ROWB;
#pragma unroll
        for (int s = 0; s < NS; ++s)
          bulk_g2s(dst + (size_t)s * SEGS * ROWB, (const unsigned char*)epi.base(s) + (size_t)row * ROWB, ROWB, &bar[b]);
      }
      q_out = q;
      nseg_out = nseg;
    };

    int p = P0, u = u0, k = 0, q = 0, nseg = 0;
    bool open = false;  // segment u has a partial in carry_s[(k & 1)] from the previous stage
    build(0, p, u, q, nseg);
    while (true) {
      const int b = k & 1;
      const int* sg = segs_s + b * (SEGS + 2);
      const bool last_cont = sg[nseg] > q;  // the stage's last segment goes on beyond q
      const int p2 = q, u2 = last_cont ? u + nseg - 1 : u + nseg;
      const bool more = p2 < P1;
      int q2 = 0, nseg2 = 0;
      if (more) build(b ^ 1, p2, u2, q2, nseg2);
      mbar_wait(&bar[b], (unsigned)(k >> 1) & 1u);

      // ---- reduce stage k out of shared memory
      const unsigned char* gs = smem + (size_t)b * stage_bytes;
      const unsigned char* ps = (smem + (size_t)b * stage_bytes) + (size_t)ROWS * ROWB_G;
      for (int i = gk; i < nseg; i += NG) {
        const int sa = sg[i], se = sg[i + 1];
        const int a = max(sa, p), e = min(se, q);
        const bool fin = se <= q;
        const bool is_head = head && (u + i == u0);
        const bool single = (se - sa == 1);
        double acc[kC][4];
        typename R::V one[kC];
        if (single) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) one[c] = R::ld(gs + (size_t)(a - p) * ROWB_G, col);
          }
        } else {
          if (i == 0 && open) {
#pragma unroll
            for (int c = 0; c < kC; ++c) {
              const int col = gl + c * kG;
              if (col < D4) {
                const double* cs = carry_s + (size_t)b * D + col * 4;
                acc[c][0] = cs[0]; acc[c][1] = cs[1]; acc[c][2] = cs[2]; acc[c][3] = cs[3];
              }
            }
          } else {
#pragma unroll
            for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
          }
          for (int r = a - p; r < e - p; ++r) {
#pragma unroll
            for (int c = 0; c < kC; ++c) {
              const int col = gl + c * kG;
              if (col < D4) R::add(acc[c], R::ld(gs + (size_t)r * ROWB_G, col));
            }
          }
        }
        if (fin && !is_head) {
          const long long row = epi.row_of(&L, P.uniq, u + i);
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) {
              float4 g;
              double s[4];
              if (single) {
                g = R::f32(one[c]);
                if (Epi::kOutF64) { s[0] = s[1] = s[2] = s[3] = 0.0; R::add(s, one[c]); }
              } else {
                g = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                                __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
                s[0] = acc[c][0]; s[1] = acc[c][1]; s[2] = acc[c][2]; s[3] = acc[c][3];
              }
              float4 pv = make_float4(0, 0, 0, 0), a0 = pv, a1 = pv;
              if (NS > 0) pv = *(const float4*)(ps + (size_t)i * ROWB + col * 16);
              if (NS > 1) a0 = *(const float4*)(ps + (size_t)SEGS * ROWB + (size_t)i * ROWB + col * 16);
              if (NS > 2) a1 = *(const float4*)(ps + (size_t)2 * SEGS * ROWB + (size_t)i * ROWB + col * 16);
              epi.apply(row, D4, col, g, s, pv, a0, a1);
            }
          }
          if (gl == 0) epi.touch(row);
        } else {
          // unfinished, or the head segment: a float64 partial goes to the next stage (shared
          // memory) or, at the end of this CTA's slice / for the head segment, to global scratch
          if (single) {
#pragma unroll
            for (int c = 0; c < kC; ++c) {
              acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
              const int col = gl + c * kG;
              if (col < D4) R::add(acc[c], one[c]);
            }
          }
          double* dst;
          if (!fin && q < P1) dst = carry_s + (size_t)(b ^ 1) * D;
          else if (is_head) dst = P.carry_h + (size_t)blockIdx.x * D;
          else dst = P.carry_t + (size_t)blockIdx.x * D;
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) {
              dst[col * 4 + 0] = acc[c][0]; dst[col * 4 + 1] = acc[c][1];
              dst[col * 4 + 2] = acc[c][2]; dst[col * 4 + 3] = acc[c][3];
            }
          }
          if (gl == 0 && !(!fin && q < P1)) {
            int32_t* f = P.flags + (size_t)blockIdx.x * 4;
            if (is_head) { f[0] = 1; f[1] = fin ? 1 : 0; }
            else { f[2] = 1; f[3] = u + i; }
          }
        }
      }
      __syncthreads();
      if (!more) break;
      open = last_cont;
      p = p2; u = u2; q = q2; nseg = nseg2; ++k;
    }
  }

  // ---- last CTA to finish stitches the segments that crossed CTA boundaries
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(P.ticket, 1u) == gridDim.x - 1) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int G = (int)gridDim.x;
  for (int c0 = gk; c0 < G; c0 += NG) {
    const volatile int32_t* f = P.flags + (size_t)c0 * 4;
    if (!f[2]) continue;
    const int useg = f[3];
    double s[kC][4];
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int col = gl + c * kG;
      if (col < D4) {
        const double* t = P.carry_t + (size_t)c0 * D + col * 4;  // __ldcg: written by other SMs
        s[c][0] = __ldcg(t); s[c][1] = __ldcg(t + 1); s[c][2] = __ldcg(t + 2); s[c][3] = __ldcg(t + 3);
      }
    }
    for (int j = c0 + 1; j < G; ++j) {
      const volatile int32_t* fj = P.flags + (size_t)j * 4;
      if (!fj[0]) break;  // cannot happen for a well-formed input; stops at an empty slice
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        const int col = gl + c * kG;
        if (col < D4) {
          const double* h = P.carry_h + (size_t)j * D + col * 4;
          s[c][0] += __ldcg(h); s[c][1] += __ldcg(h + 1); s[c][2] += __ldcg(h + 2); s[c][3] += __ldcg(h + 3);
        }
      }
      if (fj[1]) break;
    }
    const long long row = epi.row_of(&L, P.uniq, useg);
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int col = gl + c * kG;
      if (col < D4) {
        float4 pv = make_float4(0, 0, 0, 0), a0 = pv, a1 = pv;
        epi.load(row, D4, col, &pv, &a0, &a1);
        float4 g = make_float4(__double2float_rn(s[c][0]), __double2float_rn(s[c][1]),
                               __double2float_rn(s[c][2]), __double2float_rn(s[c][3]));
        epi.apply(row, D4, col, g, s[c], pv, a0, a1);
      }
    }
    if (gl == 0) epi.touch(row);
  }
}

// ------------------------------------------------------------------ host side
constexpr int kMaxGrid = EDL_SM_COUNT * 4;

struct SegWs { double* carry_h; double* carry_t; int32_t* flags; unsigned* ticket; size_t zero_bytes; size_t bytes; };
SegWs carve_seg(void* ws, int dim) {
  SegWs r; char* p = (char*)ws;
  size_t off = 0;
  r.flags = (int32_t*)(p + off); off += (size_t)kMaxGrid * 4 * sizeof(int32_t);
  r.ticket = (unsigned*)(p + off); off += 256;
  r.zero_bytes = off;  // flags + ticket are cleared before every launch
  r.carry_h = (double*)(p + off); off += (size_t)kMaxGrid * dim * sizeof(double);
  r.carry_t = (double*)(p + off); off += (size_t)kMaxGrid * dim * sizeof(double);
  r.bytes = edl_align_up(off, 256);
  return r;
}

struct StageCfg { int rows, segs, ctas_per_sm; size_t smem; };

// Stage geometry from the shared-memory budget: gradient rows + (param + slot) rows per stage,
// two stages, <= 227 KB per SM.
template <typename GT>
StageCfg stage_cfg(int dim, int n_slots) {
  const size_t rowb = (size_t)dim * 4, rowb_g = (size_t)dim * sizeof(GT);
  StageCfg best{0, 0, 1, 0};
  for (int ctas = 2; ctas >= 1; --ctas) {
    const size_t budget = (size_t)(ctas == 2 ? 110 : 220) * 1024;
    int rows = kThreads;
    while (rows >= 16) {
      int segs = n_slots > 0 ? std::max(rows / 2, 8) : rows;
      size_t stage = (size_t)rows * rowb_g + (size_t)n_slots * segs * rowb;
      size_t total = 2 * stage + 2 * (size_t)dim * 8 + 2 * (size_t)(segs + 2) * 4 + 128;
      if (total <= budget) { best = StageCfg{rows, segs, ctas, total}; break; }
      rows -= 16;
    }
    if (best.rows >= 64 || (ctas == 1 && best.rows > 0)) return best;
  }
  return best;
}

template <int kG, int kC, typename GT, typename Epi>
int launch_tile(const void* grad, int D4, const int32_t* order, const int32_t* seg,
                const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
                const int64_t* layout, Epi epi, const SegWs& W, cudaStream_t st) {
  (void)max_unique;
  const StageCfg cfg = stage_cfg<GT>(D4 * 4, Epi::kSlots);
  EDL_REQUIRE(cfg.rows > 0, EDL_ELIMIT, "segreduce: dim=%d rows do not fit a shared-memory stage", D4 * 4);
  auto kern = k_seg_tile<kG, kC, GT, Epi>;
  EDL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
  EDL_CUDA(cudaMemsetAsync(W.flags, 0, W.zero_bytes, st));
  TileParams P;
  P.grad = (const unsigned char*)grad; P.D4 = D4; P.order = order; P.seg = seg; P.n_unique = n_unique;
  P.uniq = uniq; P.layout = layout; P.rows_per_stage = cfg.rows; P.segs_per_stage = cfg.segs;
  P.carry_h = W.carry_h; P.carry_t = W.carry_t; P.flags = W.flags; P.ticket = W.ticket;
  const int grid = EDL_SM_COUNT * cfg.ctas_per_sm;
  kern<<<grid, kThreads, cfg.smem, st>>>(P, epi);
  EDL_LAUNCH_CHECK("k_seg_tile");
  return EDL_OK;
}

template <typename GT, typename Epi>
int dispatch_dim(const void* grad, int dim, const int32_t* order, const int32_t* seg,
                 const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
                 const int64_t* layout, Epi epi, const SegWs& W, cudaStream_t st) {
  const int D4 = dim / 4;
#define EDL_GO(G, C) return launch_tile<G, C, GT, Epi>(grad, D4, order, seg, n_unique, max_unique, uniq, layout, epi, W, st)
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

int check_common(const char* who, const void* grad, int dim, const int32_t* order,
                 const int32_t* seg, const int32_t* n_unique, int64_t max_unique, void* ws,
                 size_t ws_bytes, SegWs* W) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "%s: dim=%d must be a positive multiple of 4", who, dim);
  EDL_REQUIRE(max_unique >= 0, EDL_EINVAL, "%s: max_unique negative", who);
  EDL_REQUIRE(grad && order && seg && n_unique, EDL_EINVAL, "%s: null pointer", who);
  EDL_REQUIRE((((uintptr_t)grad) & 15) == 0, EDL_EINVAL, "%s: grad must be 16-byte aligned", who);
  *W = carve_seg(ws, dim);
  EDL_REQUIRE(ws && ws_bytes >= W->bytes, EDL_EWORKSPACE, "%s: workspace %zu < required %zu", who,
              ws_bytes, W->bytes);
  return EDL_OK;
}

}  // namespace

extern "C" size_t edl_segreduce_workspace_bytes(int64_t max_unique, int dim) {
  (void)max_unique;
  return carve_seg(nullptr, dim > 0 ? dim : 4).bytes;
}

extern "C" int edl_segreduce(const float* grad, int dim, const int32_t* order,
                             const int32_t* seg_offsets, const int32_t* n_unique_dev,
                             int64_t max_unique, const int32_t* out_pos, void* g_out, int out_f64,
                             void* workspace, size_t workspace_bytes, edl_stream_t stream) {
  if (max_unique == 0) return EDL_OK;
  SegWs W;
  int rc = check_common("edl_segreduce", grad, dim, order, seg_offsets, n_unique_dev, max_unique,
                        workspace, workspace_bytes, &W);
  if (rc) return rc;
  EDL_REQUIRE(g_out && (((uintptr_t)g_out) & 15) == 0, EDL_EINVAL, "edl_segreduce: g_out null or misaligned");
  cudaStream_t st = (cudaStream_t)stream;
  if (out_f64) {
    EpiWrite64 epi{(double*)g_out, out_pos};
    return dispatch_dim<float, EpiWrite64>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, nullptr,
                                           nullptr, epi, W, st);
  }
  EpiWrite epi{(float4*)g_out, out_pos};
  return dispatch_dim<float, EpiWrite>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, nullptr,
                                       nullptr, epi, W, st);
}

extern "C" int edl_segreduce_update(float* slab, float* slot0, float* slot1, int dim,
                                    const int64_t* uniq, const int32_t* order,
                                    const int32_t* seg_offsets, const int32_t* n_unique_dev,
                                    int64_t max_unique, const void* grad, int grad_f64,
                                    const int64_t* layout, const edl_opt_params* opt,
                                    uint32_t* touched, void* workspace, size_t workspace_bytes,
                                    edl_stream_t stream) {
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
    EpiAdagrad epi{(float4*)slab, (float4*)slot0, touched, o};
    return grad_f64 ? dispatch_dim<double, EpiAdagrad>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st)
                    : dispatch_dim<float, EpiAdagrad>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st);
  }
  if (opt->kind == EDL_OPT_ADAM) {
    EDL_REQUIRE(slot1, EDL_EINVAL, "edl_segreduce_update: Adam needs slot1 (v)");
    EpiAdam epi{(float4*)slab, (float4*)slot0, (float4*)slot1, touched, o};
    return grad_f64 ? dispatch_dim<double, EpiAdam>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st)
                    : dispatch_dim<float, EpiAdam>(grad, dim, order, seg_offsets, n_unique_dev, max_unique, uniq, layout, epi, W, st);
  }
  edl_set_error("edl_segreduce_update: unknown optimizer kind %d", opt->kind);
  return EDL_EINVAL;
}
