// k_seg_tile: shared-memory staged, position-balanced segmented reduce + optimizer -- the alternative
// kernel family (EDL_SEG_IMPL=tile).  Persistent CTAs over slices of the SORTED positions, rows copied
// asynchronously (LDGSTS; TMA bulk with -DEDL_TILE_BULK) into a 3-stage shared-memory ring, index
// metadata prefetched one window ahead, float64 carries across stages and CTAs.  Own translation unit.
#include "seg_common.cuh"

namespace {

// ------------------------------------------------------------------ the tile kernel
struct TileParams {
  const unsigned char* grad;  // [N, D] rows of GT
  int D4;
  const int32_t* order;
  const int32_t* seg;
  const int32_t* n_unique;
  const int64_t* uniq;
  const int64_t* layout;
  int rows_per_stage;  // ROWS
  int segs_per_stage;  // SEGS <= 64
  double* carry_h;     // [grid][D]  partial of the segment that continues from the previous CTA
  double* carry_t;     // [grid][D]  partial of the last segment when it continues into the next CTA
  int32_t* flags;      // [grid][4]  h_valid, h_ends, t_valid, t_seg
  unsigned* ticket;
  unsigned part_offset; // byte offset of the long-run partial buffer in dynamic shared memory
  int debug;            // 0 normal; 1 skip the reduce; 2 skip the copies (timing experiments only)
};

constexpr int kLongRun = 12; // in-stage runs longer than this are summed by all lane groups together
constexpr int kStages = 3;   // stage buffers: two stages of copies in flight while one is reduced
constexpr int kMP = 256;     // positions per metadata window (== kThreads: one element per thread)
constexpr int kMS = kMP + 2; // segment offsets per window: every segment starting in it + the end

template <int kG, int kC, typename GT, typename Epi, bool kBulk>
__global__ void __launch_bounds__(kThreads)
k_seg_tile(TileParams P, Epi epi) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ EdlLayoutSmem L;
  __shared__ int s_last;
  __shared__ int desc_s[kStages][8];   // p, q, u, nseg, open
  __shared__ uint64_t bar[kStages];    // one mbarrier per stage buffer (TMA bulk-copy completion)
  constexpr int NG = kThreads / kG;
  constexpr int NS = Epi::kSlots;
  typedef Row<GT> R;
  const int tid = threadIdx.x, lane = tid & 31;
  const int gk = tid / kG, gl = tid % kG;
  const int D4 = P.D4, D = D4 * 4;
  const int ROWS = P.rows_per_stage, SEGS = P.segs_per_stage;
  const unsigned ROWB = (unsigned)D4 * 16u;              // fp32 row bytes
  const unsigned ROWB_G = (unsigned)D4 * R::kBytes4;     // gradient row bytes
  const size_t stage_bytes = (size_t)ROWS * ROWB_G + (size_t)NS * SEGS * ROWB;
  const int chg_n = (int)(ROWB_G / 16), ch_n = (int)(ROWB / 16);
  const int chg_sh = (chg_n & (chg_n - 1)) == 0 ? (__ffs(chg_n) - 1) : -1;
  const int ch_sh = (ch_n & (ch_n - 1)) == 0 ? (__ffs(ch_n) - 1) : -1;
  // shared-memory map: [kStages stage buffers][carry 2 x D f64][uniq windows][order][seg][stage segs]
  double* carry_s = (double*)(smem + kStages * stage_bytes);
  long long* wuq_s = (long long*)(carry_s + 2 * D);            // [2][kMS]
  int* word_s = (int*)(wuq_s + 2 * kMS);                       // [2][kMP]
  int* wseg_s = word_s + 2 * kMP;                              // [2][kMS]
  int* segs_s = wseg_s + 2 * kMS;                              // [kStages][SEGS + 1]
  double* part_s = (double*)(smem + P.part_offset);            // [NG][D] long-run partial sums
  long long* srow_s = (long long*)(part_s + (size_t)NG * D);   // [kStages][SEGS] slab row of each segment

  edl_layout_load(&L, P.layout);
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < kStages; ++i) mbar_init(&bar[i], 1);
    fence_proxy_async();
  }
  __syncthreads();

  const int U = *P.n_unique;
  const int N = U > 0 ? P.seg[U] : 0;
  int per = (int)(((long long)N + gridDim.x - 1) / gridDim.x);
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

    // ---- metadata windows: order / seg / uniq slices for kMP positions, prefetched into
    // registers one window ahead so that building a stage never waits on global memory
    int w_pos = P0, w_seg = u0;          // window being built from: first position, first segment
    int wb = 0;                          // its buffer (0/1)
    int pf_ord = 0, pf_seg0 = 0, pf_seg1 = 0;
    long long pf_uq0 = 0, pf_uq1 = 0;
    int pf_pos = 0, pf_segbase = 0;      // what the prefetch registers hold
    auto win_fetch = [&](int pos, int segbase) {   // issue the loads (no use yet)
      pf_pos = pos; pf_segbase = segbase;
      pf_ord = (pos + tid < N) ? P.order[pos + tid] : 0;
      pf_seg0 = P.seg[min(segbase + tid, U)];
      pf_seg1 = P.seg[min(segbase + kThreads + tid, U)];
      pf_uq0 = P.uniq ? P.uniq[min(segbase + tid, U - 1)] : 0;
      pf_uq1 = P.uniq ? P.uniq[min(segbase + kThreads + tid, U - 1)] : 0;
    };
    auto win_store = [&](int b) {                  // registers -> shared (caller syncs)
      word_s[b * kMP + tid] = pf_ord;
      wseg_s[b * kMS + tid] = pf_seg0;
      wuq_s[b * kMS + tid] = pf_uq0;
      if (kThreads + tid < kMS) {
        wseg_s[b * kMS + kThreads + tid] = pf_seg1;
        wuq_s[b * kMS + kThreads + tid] = pf_uq1;
      }
    };
    auto win_next = [&](int b, int pos, int segbase, int* npos, int* nseg) {
      // first segment of the window that starts at pos + kMP: largest i with seg[i] <= pos + kMP
      const int* ws = wseg_s + b * kMS;
      const int target = pos + kMP;
      int l = 0, h = kMS;   // ws[0] <= pos < target; entries beyond U are clamped to N >= target? (N may equal target)
      while (h - l > 1) {
        int mid = (l + h) >> 1;
        if (ws[mid] <= target && segbase + mid < U) l = mid; else h = mid;
      }
      *npos = target;
      *nseg = segbase + l;
    };
    win_fetch(w_pos, w_seg);
    win_store(0);
    __syncthreads();
    {
      int np, ns;
      win_next(0, w_pos, w_seg, &np, &ns);
      if (np < P1) win_fetch(np, ns);
    }

    // ---- stage builder: extent from the window's segment offsets, then the async copies
    int bp = P0, bu = u0;        // build cursor: next position / segment to stage
    bool bopen = false;
    auto build = [&](int st) {
      if (bp >= w_pos + kMP) {   // cursor left the window: bring in the prefetched one
        wb ^= 1;
        win_store(wb);
        w_pos = pf_pos; w_seg = pf_segbase;
        __syncthreads();
        int np, ns;
        win_next(wb, w_pos, w_seg, &np, &ns);
        if (np < P1) win_fetch(np, ns);
      }
      const int* ws = wseg_s + wb * kMS;
      const int so = bu - w_seg;                       // cursor segment inside the window
      auto seg_at = [&](int i) -> int {                // offset of segment bu + i
        const int j = so + i;
        return (j < kMS && bu + i <= U) ? ws[j] : 0x7fffffff;
      };
      const int lim = min(min(bp + ROWS, w_pos + kMP), P1);
      const int q = min(lim, seg_at(SEGS));
      int nseg = 0;
      for (int i0 = 0; i0 < SEGS; i0 += 32)
        nseg += __popc(__ballot_sync(0xffffffffu, (i0 + lane < SEGS) && seg_at(i0 + lane) < q));
      int* sg = segs_s + st * (SEGS + 1);
      if (tid <= nseg) sg[tid] = seg_at(tid);
      // segments whose run inside this stage is long (a hot id): bit i of a 64-bit mask
      auto long_of = [&](int i) -> bool {
        return i < nseg && (min(seg_at(i + 1), q) - max(seg_at(i), bp)) > kLongRun;
      };
      const unsigned lm0 = __ballot_sync(0xffffffffu, long_of(lane));
      const unsigned lm1 = SEGS > 32 ? __ballot_sync(0xffffffffu, long_of(lane + 32)) : 0u;
      if (tid == 0) {
        desc_s[st][0] = bp; desc_s[st][1] = q; desc_s[st][2] = bu; desc_s[st][3] = nseg;
        desc_s[st][4] = bopen ? 1 : 0; desc_s[st][5] = (int)lm0; desc_s[st][6] = (int)lm1;
      }
      unsigned char* gdst = smem + (size_t)st * stage_bytes;
      unsigned char* pdst = gdst + (size_t)ROWS * ROWB_G;
      // producer: one TMA bulk copy per row (256 B at dim 64), bytes counted on the stage's
      // mbarrier.  One instruction per row instead of 16 LDGSTS per row keeps the issue slots
      // for the float64 adds and the IEEE sqrt / divide of the optimizer epilogue.
      auto fin_of = [&](int i) -> bool {   // segment bu + i ends inside this stage (and is not the head)
        return i < nseg && seg_at(i + 1) <= q && !(head && bu + i == u0);
      };
      bool fin = false;
      int nfin = 0;
      if (NS > 0) {
        fin = fin_of(tid);
        nfin = __popc(__ballot_sync(0xffffffffu, fin_of(lane)));
        if (SEGS > 32) nfin += __popc(__ballot_sync(0xffffffffu, fin_of(lane + 32)));
      }
      long long my_row = 0;
      if (tid < nseg) {
        if (NS > 0) my_row = edl_key_to_row(&L, wuq_s[wb * kMS + so + tid]);
        else my_row = epi.row_of(&L, P.uniq, bu + tid);
        srow_s[st * SEGS + tid] = my_row;
      }
      if (P.debug == 2) {
        if (kBulk && tid == 0) mbar_expect_tx(&bar[st], 0u);
      } else if (kBulk) {
        if (tid == 0) mbar_expect_tx(&bar[st], (unsigned)(q - bp) * ROWB_G + (unsigned)nfin * ROWB * NS);
        if (tid < q - bp) {
          const long long j = word_s[wb * kMP + (bp - w_pos) + tid];
          bulk_g2s(gdst + (size_t)tid * ROWB_G, P.grad + (size_t)j * ROWB_G, ROWB_G, &bar[st]);
        }
        if (NS > 0 && fin) {
          const long long row = my_row;
#pragma unroll
          for (int s = 0; s < NS; ++s)
            bulk_g2s(pdst + (size_t)s * SEGS * ROWB + (size_t)tid * ROWB,
                     (const unsigned char*)epi.base(s) + (size_t)row * ROWB, ROWB, &bar[st]);
        }
      } else {
        const int* wo = word_s + wb * kMP + (bp - w_pos);
        const int chg = (int)(ROWB_G / 16);
        const int total_g = (q - bp) * chg;
        if (chg_sh >= 0) {   // power-of-two row width: shifts instead of a division
          for (int idx = tid; idx < total_g; idx += kThreads) {
            const int r = idx >> chg_sh;
            const long long j = wo[r];
            cp_async16(gdst + (size_t)idx * 16, P.grad + (size_t)j * ROWB_G + (size_t)(idx & (chg - 1)) * 16);
          }
        } else {
          for (int idx = tid; idx < total_g; idx += kThreads) {
            const int r = idx / chg, c = idx - r * chg;
            const long long j = wo[r];
            cp_async16(gdst + (size_t)r * ROWB_G + c * 16, P.grad + (size_t)j * ROWB_G + c * 16);
          }
        }
        if (NS > 0) {
          const long long* wu = wuq_s + wb * kMS + so;
          (void)wu;
          const int ch = (int)(ROWB / 16);
          const int total_p = nseg * ch;
          // finishing segments as a bit mask, slab rows from the values thread i just computed
          const unsigned long long fmask = (unsigned long long)__ballot_sync(0xffffffffu, fin_of(lane)) |
              (SEGS > 32 ? ((unsigned long long)__ballot_sync(0xffffffffu, fin_of(lane + 32)) << 32) : 0ull);
          __syncthreads();   // srow_s of this stage visible
          for (int idx = tid; idx < total_p; idx += kThreads) {
            const int i = ch_sh >= 0 ? (idx >> ch_sh) : (idx / ch);
            const int c = idx - i * ch;
            if ((fmask >> i) & 1ull) {
              const long long row = srow_s[st * SEGS + i];
#pragma unroll
              for (int s = 0; s < NS; ++s)
                cp_async16(pdst + (size_t)s * SEGS * ROWB + (size_t)idx * 16,
                           (const unsigned char*)epi.base(s) + (size_t)row * ROWB + c * 16);
            }
          }
        }
      }
      const bool last_cont = seg_at(nseg) > q;
      bu = last_cont ? bu + nseg - 1 : bu + nseg;
      bopen = last_cont;
      bp = q;
    };

    // ---- pipeline: stages k+1, k+2 in flight while stage k is reduced
    if (bp < P1) build(0);
    if (!kBulk) cp_async_commit();
    if (bp < P1) build(1);
    if (!kBulk) cp_async_commit();
    __syncthreads();   // stage descriptors of the first two stages visible
    for (int k = 0;; ++k) {
      const int st = k % kStages;
      if (bp < P1) build((k + 2) % kStages);
      if (kBulk) {
        mbar_wait(&bar[st], (unsigned)(k / kStages) & 1u);
      } else {
        cp_async_commit();
        cp_async_wait<2>();
        __syncthreads();
      }

      const int p = desc_s[st][0], q = desc_s[st][1], u = desc_s[st][2], nseg = desc_s[st][3];
      const bool open = desc_s[st][4] != 0;
      const int b = k & 1;  // carry buffer read by this stage; it writes (b ^ 1)
      const int* sg = segs_s + st * (SEGS + 1);
      const unsigned char* gs = smem + (size_t)st * stage_bytes;
      const unsigned char* ps = gs + (size_t)ROWS * ROWB_G;
      // finishes one segment (or one in-stage run of it): epilogue, or a float64 partial to the
      // next stage / to the CTA-boundary scratch
      auto finish = [&](int i, bool fin, bool is_head, bool single, double (&acc)[kC][4],
                        typename R::V (&one)[kC]) {
        if (fin && !is_head) {
          const long long row = srow_s[st * SEGS + i];
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
      };

      // pass 1: one lane group per segment, runs of up to kLongRun rows
      bool any_long = false;
      const int nseg_r = P.debug == 1 ? 0 : nseg;
      for (int i = gk; i < nseg_r; i += NG) {
        const int sa = sg[i], se = sg[i + 1];
        const int a = max(sa, p), e = min(se, q);
        if (e - a > kLongRun) continue;
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
#pragma unroll 4
          for (int r = a - p; r < e - p; ++r) {
#pragma unroll
            for (int c = 0; c < kC; ++c) {
              const int col = gl + c * kG;
              if (col < D4) R::add(acc[c], R::ld(gs + (size_t)r * ROWB_G, col));
            }
          }
        }
        finish(i, fin, is_head, single, acc, one);
      }
      // pass 2: long in-stage runs (a hot id): all lane groups sum strided rows in float64, the
      // partial rows are added in group order by group 0.  Uniform control flow (block barriers).
      unsigned long long lmask = P.debug == 1 ? 0ull
          : ((unsigned long long)(unsigned)desc_s[st][5] | ((unsigned long long)(unsigned)desc_s[st][6] << 32));
      while (lmask) {
        const int i = __ffsll((long long)lmask) - 1;
        lmask &= lmask - 1;
        const int sa = sg[i], se = sg[i + 1];
        const int a = max(sa, p), e = min(se, q);
        any_long = true;
        double acc[kC][4];
        typename R::V one[kC];
#pragma unroll
        for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
#pragma unroll 2
        for (int r = a - p + gk; r < e - p; r += NG) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) R::add(acc[c], R::ld(gs + (size_t)r * ROWB_G, col));
          }
        }
#pragma unroll
        for (int c = 0; c < kC; ++c) {
          const int col = gl + c * kG;
          if (col < D4) {
            double* dst = part_s + (size_t)gk * D + col * 4;
            dst[0] = acc[c][0]; dst[1] = acc[c][1]; dst[2] = acc[c][2]; dst[3] = acc[c][3];
          }
        }
        __syncthreads();
        if (gk == 0) {
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
          for (int g2 = 0; g2 < NG; ++g2) {
#pragma unroll
            for (int c = 0; c < kC; ++c) {
              const int col = gl + c * kG;
              if (col < D4) {
                const double* src = part_s + (size_t)g2 * D + col * 4;
                acc[c][0] += src[0]; acc[c][1] += src[1]; acc[c][2] += src[2]; acc[c][3] += src[3];
              }
            }
          }
          finish(i, se <= q, head && (u + i == u0), false, acc, one);
        }
        __syncthreads();
      }
      (void)any_long;
      __syncthreads();
      if (q >= P1) break;
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
struct StageCfg { int rows, segs, ctas_per_sm; size_t smem; };

// bytes before the long-run partial buffer ([NG][dim] float64 = kThreads * 4 * 8 bytes per
// 128 columns of lane-group width)
size_t tile_part_offset(int dim, size_t rowb_g, int n_slots, int rows, int segs) {
  const size_t rowb = (size_t)dim * 4;
  const size_t stage = (size_t)rows * rowb_g + (size_t)n_slots * segs * rowb;
  size_t off = kStages * stage + 2 * (size_t)dim * 8 + 2 * (size_t)kMS * 8 + 2 * (size_t)kMP * 4 +
               2 * (size_t)kMS * 4 + (size_t)kStages * (segs + 1) * 4;
  return edl_align_up(off, 16);
}

size_t tile_smem_bytes(int dim, size_t rowb_g, int n_slots, int rows, int segs) {
  const size_t rowb = (size_t)dim * 4;
  const size_t stage = (size_t)rows * rowb_g + (size_t)n_slots * segs * rowb;
  return tile_part_offset(dim, rowb_g, n_slots, rows, segs) + (size_t)kThreads * 4 * 8 * ((dim + 127) / 128) +
         (size_t)kStages * segs * 8 + 128;
}

// Stage geometry from the shared-memory budget: three stage buffers of gradient rows +
// (param + slot) rows, two CTAs per SM when a stage of >= 32 rows fits, else one.
template <typename GT>
StageCfg stage_cfg(int dim, int n_slots) {
  const size_t rowb_g = (size_t)dim * sizeof(GT);
  for (int ctas = 2; ctas >= 1; --ctas) {
    const size_t budget = (size_t)(ctas == 2 ? 111 : 224) * 1024;
    for (int rows = 128; rows >= 8; rows -= 8) {
      const int segs = n_slots > 0 ? std::min(64, std::max(rows / 2, 4)) : std::min(64, rows);
      const size_t total = tile_smem_bytes(dim, rowb_g, n_slots, rows, segs);
      if (total <= budget) {
        if (rows >= 32 || ctas == 1) return StageCfg{rows, segs, ctas, total};
        break;
      }
    }
  }
  return StageCfg{0, 0, 1, 0};
}

template <int kG, int kC, typename GT, typename Epi>
int launch_tile(const void* grad, int D4, const int32_t* order, const int32_t* seg,
                const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
                const int64_t* layout, Epi epi, const SegWs& W, cudaStream_t st) {
  (void)max_unique;
  const StageCfg cfg = stage_cfg<GT>(D4 * 4, Epi::kSlots);
  EDL_REQUIRE(cfg.rows > 0, EDL_ELIMIT, "segreduce: dim=%d rows do not fit a shared-memory stage", D4 * 4);
  static int s_bulk = -1, s_debug = 0;
  if (s_bulk < 0) {
    const char* e = getenv("EDL_SEG_COPY");   // "ldgsts" (default) or "bulk": producer copy engine
    s_bulk = (e && e[0] == 'b') ? 1 : 0;
    const char* d = getenv("EDL_SEG_DEBUG");
    s_debug = d ? atoi(d) : 0;
  }
#ifdef EDL_TILE_BULK
  auto kern = s_bulk ? k_seg_tile<kG, kC, GT, Epi, true> : k_seg_tile<kG, kC, GT, Epi, false>;
#else
  (void)s_bulk;
  auto kern = k_seg_tile<kG, kC, GT, Epi, false>;
#endif
  static int attr_smem[64] = {0};  // per device: opt in to > 48 KB dynamic shared memory once
  int dev = 0;
  EDL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || attr_smem[dev] < (int)cfg.smem) {
#ifdef EDL_TILE_BULK
    EDL_CUDA(cudaFuncSetAttribute(k_seg_tile<kG, kC, GT, Epi, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
#endif
    EDL_CUDA(cudaFuncSetAttribute(k_seg_tile<kG, kC, GT, Epi, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.smem));
    if (dev >= 0 && dev < 64) attr_smem[dev] = (int)cfg.smem;
  }
  TileParams P;
  P.grad = (const unsigned char*)grad; P.D4 = D4; P.order = order; P.seg = seg; P.n_unique = n_unique;
  P.uniq = uniq; P.layout = layout; P.rows_per_stage = cfg.rows; P.segs_per_stage = cfg.segs;
  P.carry_h = W.carry_h; P.carry_t = W.carry_t; P.flags = W.flags; P.ticket = W.ticket;
  P.debug = s_debug;
  P.part_offset = (unsigned)tile_part_offset(D4 * 4, (size_t)D4 * 4 * sizeof(GT), Epi::kSlots, cfg.rows, cfg.segs);
  const int grid = EDL_SM_COUNT * cfg.ctas_per_sm;
  kern<<<grid, kThreads, cfg.smem, st>>>(P, epi);
  EDL_LAUNCH_CHECK("k_seg_tile");
  return EDL_OK;
}

}  // namespace

// type-erased entry of this kernel family (declared in update.cu)
int edl_seg_run_tile(const SegCall& c, const SegWs& W) {
  return seg_with_epi(c, [&](auto gt, auto epi) {
    using GT = decltype(gt);
    using Epi = decltype(epi);
    return seg_with_dim(c.dim, [&](auto G, auto C) {
      return launch_tile<decltype(G)::value, decltype(C)::value, GT, Epi>(
          c.grad, c.dim / 4, c.order, c.seg, c.n_unique, c.max_unique, c.uniq, c.layout, epi, W, c.st);
    });
  });
}
