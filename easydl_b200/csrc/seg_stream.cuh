// Streamed segmented reduce + optimizer (the default main kernel of A4/A5): persistent, warp-autonomous.
//
// Every warp pulls tickets of 32 consecutive unique ids from a device counter (dynamic balance under
// any skew) and walks them with one lane group (dim/4 lanes, as in seg_direct.cuh) per id.  The rows
// an id needs -- its gradient rows one by one, and with the last of them its parameter / slot rows --
// are copied asynchronously (cp.async, LDGSTS) into a ring of kNS stages that is private to the warp,
// kNS - 1 stages ahead of the arithmetic:
//
//   iteration t:   issue(slot t + kNS - 1)  |  cp.async.wait_group kNS - 1  |  consume(slot t)
//
// so the loads in flight do not live in registers (the register path spilled at 32 registers and its
// spill stores waited for the very loads they were meant to overlap), a warp's loads are in flight
// all the time instead of once per id, and the index work (segment bounds, first row, key -> slab row)
// is done once per id by the lane that holds the id's record instead of by every lane of the group.
// A lane only ever reads back the 16-byte pieces it copied itself, and keeps the slot's descriptor in
// its own shared-memory word, so the loop needs no barrier: only the shuffles that hand out ids.
// Ids with more than kLongThr gradient rows are skipped: k_seg_classify lists them for the chunk
// kernels (seg_direct.cuh), which run beside this kernel on the helper stream.
// Summation order inside an id is its row order whatever the schedule: results are deterministic.
#pragma once
#include <atomic>

#include "seg_common.cuh"

namespace {

template <int kG, int kC, typename GT, typename Epi>
struct StreamShape {
  typedef Row<GT> R;
  static constexpr int kWarps = kThreads / 32;
  static constexpr int kNG = 32 / kG;                        // lane groups (ids in flight) per warp
  static constexpr int kH = R::kBytes4 / 16;                 // 16-byte pieces per gradient column of a lane
  static constexpr int kSlotP = Epi::kSlots * kC;            // parameter / slot pieces per lane
  static constexpr int kNP = kSlotP + kC * kH;               // pieces per lane and stage
  static constexpr int kStageBytes = kNP * 512 + 256;        // piece q of lane l at q * 512 + l * 16; then 8 B / lane
  static constexpr int kBudget = 200 * 1024;
  static constexpr int kNS = kWarps * 4 * kStageBytes <= kBudget ? 4 : (kWarps * 3 * kStageBytes <= kBudget ? 3 : 2);
  static constexpr int kSmem = kWarps * kNS * kStageBytes;
  static constexpr int kMinCtas = kC == 1 ? 3 : (kC == 2 ? 2 : 1);
};

template <int kG, int kC, typename GT, typename Epi>
__global__ void __launch_bounds__(kThreads, (StreamShape<kG, kC, GT, Epi>::kMinCtas))
k_seg_stream(const unsigned char* __restrict__ grad, int D4, const int32_t* __restrict__ order,
             const int32_t* __restrict__ seg, const int32_t* __restrict__ seg_first,
             const int32_t* __restrict__ n_unique, const int64_t* __restrict__ uniq,
             const int64_t* __restrict__ layout, Epi epi, int32_t* __restrict__ ticket_counter, int kLongThr) {
  typedef StreamShape<kG, kC, GT, Epi> S;
  typedef Row<GT> R;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  epi.prepare();
  __syncthreads();
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int kNS = S::kNS, kNP = S::kNP, kSlotP = S::kSlotP, kH = S::kH;
  const int U = *n_unique;
  const unsigned lane = threadIdx.x & 31u;
  const int warp = (int)threadIdx.x >> 5;
  const int gl = (int)lane % kG, grp = (int)lane / kG;
  unsigned char* const wbase = smem_raw + (size_t)warp * (kNS * S::kStageBytes) + lane * 16;
  unsigned char* const dbase = smem_raw + (size_t)warp * (kNS * S::kStageBytes) + kNP * 512 + lane * 8;
  const int W = (int)gridDim.x * S::kWarps;
  const unsigned ROWB_G = (unsigned)D4 * R::kBytes4;
  const uint64_t pol = l2_policy_evict_first();
  bool colok[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) colok[c] = gl + c * kG < D4;

  // ---- id records: lane l holds id 32 * ticket + l.  Current ticket decoded, next ticket raw (its
  // loads were issued one ticket ago), the ticket after that still on its way back from the counter.
  int r_s0 = 0, r_cnt = 0, r_first = 0;
  long long r_row = 0;
  int n_s0 = 0, n_s1 = 0, n_first = -1;
  long long n_tok = 0;
  unsigned avail = 0;        // eligible ids of the current ticket not yet handed to a group
  bool done = false;         // the current ticket starts beyond U: no work left
  int tk_next = (int)blockIdx.x * S::kWarps + warp;   // first ticket: one per warp, no atomic
  int tk_pend = tk_next + W;                           // lane 0: second ticket, then counter values

  auto load_raw = [&](int tk) {
    const long long u = (long long)tk * 32 + lane;
    n_s0 = n_s1 = 0;
    n_first = -1;
    if (u < U) {
      n_s0 = __ldg(seg + u);
      n_s1 = __ldg(seg + u + 1);
      if (seg_first) n_first = __ldg(seg_first + u);
      n_tok = epi.token(uniq, (int)u);
    }
  };
  auto promote = [&]() {   // next ticket becomes current; fetch the one after; ask for a further one
    r_s0 = n_s0;
    r_cnt = n_s1 - n_s0;
    r_first = n_first;
    const bool mine = r_cnt > 0 && r_cnt <= kLongThr;
    r_row = mine ? epi.row_from(&L, n_tok) : 0;
    if (mine && r_first < 0) r_first = __ldg(order + r_s0);
    avail = __ballot_sync(kFull, mine);
    done = (long long)tk_next * 32 >= U;
    tk_next = __shfl_sync(kFull, tk_pend, 0);
    load_raw(tk_next);
    if (lane == 0) tk_pend = atomicAdd(ticket_counter, 1) + 2 * W;
  };
  load_raw(tk_next);
  promote();

  // ---- the lane group's current id
  int pos = 0, end = 0;      // next position in order[] and segment end; pos == end: none
  int j = 0;                 // gradient row of position pos
  long long row = 0;         // epilogue row (slab row / output position)
  bool fresh = false;        // the next slot is the id's first
  double acc[kC][4];
#pragma unroll
  for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;

  int st_i = 0, st_c = 0, empties = 0;
  for (int it = 0;; ++it) {
    // ---- groups that finished their id take the next eligible one of the ticket
#pragma unroll
    for (int g = 0; g < S::kNG; ++g) {
      const int need = __shfl_sync(kFull, (int)(pos == end), g * kG);
      if (need) {   // warp-uniform
        while (avail == 0 && !done) promote();
        if (avail != 0) {
          const int l = __ffs(avail) - 1;
          avail &= avail - 1;
          const int t_s0 = __shfl_sync(kFull, r_s0, l);
          const int t_cnt = __shfl_sync(kFull, r_cnt, l);
          const int t_first = __shfl_sync(kFull, r_first, l);
          const long long t_row = __shfl_sync(kFull, r_row, l);
          if (grp == g) { pos = t_s0; end = t_s0 + t_cnt; j = t_first; row = t_row; fresh = true; }
        }
      }
    }
    // ---- issue one slot: the gradient row at `pos`; with an id's last row also its parameter / slot rows
    unsigned long long d = 0;
    {
      unsigned char* sb = wbase + st_i * S::kStageBytes;
      if (pos < end) {
        const bool last = pos + 1 == end;
        const unsigned char* gsrc = grad + (size_t)j * ROWB_G + (size_t)gl * R::kBytes4;
#pragma unroll
        for (int c = 0; c < kC; ++c) {
          if (colok[c]) {
#pragma unroll
            for (int h = 0; h < kH; ++h)
              cp_async16_hint(sb + (kSlotP + c * kH + h) * 512, gsrc + (size_t)c * kG * R::kBytes4 + h * 16, pol);
          }
        }
        if (Epi::kSlots > 0 && last) {
#pragma unroll
          for (int s = 0; s < Epi::kSlots; ++s) {
            const unsigned char* b = (const unsigned char*)epi.base(s) + ((size_t)row * D4 + gl) * 16;
#pragma unroll
            for (int c = 0; c < kC; ++c)
              if (colok[c]) cp_async16_hint(sb + (s * kC + c) * 512, b + (size_t)c * kG * 16, pol);
          }
        }
        d = 1ull | (fresh ? 2ull : 0ull) | (last ? 4ull : 0ull) | ((unsigned long long)row << 8);
        fresh = false;
        ++pos;
        if (pos < end) j = __ldg(order + pos);   // needed one iteration from now
      }
      *(unsigned long long*)(dbase + st_i * S::kStageBytes) = d;
      cp_async_commit();
    }
    const unsigned anyv = __ballot_sync(kFull, (d & 1ull) != 0);
    st_i = st_i + 1 == kNS ? 0 : st_i + 1;
    if (it < kNS - 1) continue;   // ring not yet primed
    cp_async_wait<kNS - 1>();
    // ---- consume the oldest slot
    {
      const unsigned char* cb = wbase + st_c * S::kStageBytes;
      const unsigned long long cd = *(const unsigned long long*)(dbase + st_c * S::kStageBytes);
      if (cd & 1ull) {
        const bool first = (cd & 2ull) != 0, last = (cd & 4ull) != 0;
        const long long crow = (long long)(cd >> 8);
        typename R::V v[kC];
#pragma unroll
        for (int c = 0; c < kC; ++c) {
          if (colok[c]) {
            if constexpr (kH == 1) {
              v[c] = *(const float4*)(cb + (kSlotP + c) * 512);
            } else {
              v[c].a = *(const double2*)(cb + (kSlotP + c * 2) * 512);
              v[c].b = *(const double2*)(cb + (kSlotP + c * 2 + 1) * 512);
            }
          }
        }
        if (first && last) {   // an id that occurs once: its gradient row is the sum
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            if (colok[c]) {
              double s[4] = {0.0, 0.0, 0.0, 0.0};
              if (Epi::kOutF64) R::add(s, v[c]);
              float4 p = make_float4(0, 0, 0, 0), a0 = p, a1 = p;
              if (Epi::kSlots > 0) p = *(const float4*)(cb + (0 * kC + c) * 512);
              if (Epi::kSlots > 1) a0 = *(const float4*)(cb + (1 * kC + c) * 512);
              if (Epi::kSlots > 2) a1 = *(const float4*)(cb + (2 * kC + c) * 512);
              epi.apply(crow, D4, gl + c * kG, R::f32(v[c]), s, p, a0, a1);
            }
          }
          if (gl == 0) epi.touch(crow);
        } else {
          if (first) {
#pragma unroll
            for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
          }
#pragma unroll
          for (int c = 0; c < kC; ++c)
            if (colok[c]) R::add(acc[c], v[c]);
          if (last) {
#pragma unroll
            for (int c = 0; c < kC; ++c) {
              if (colok[c]) {
                float4 p = make_float4(0, 0, 0, 0), a0 = p, a1 = p;
                if (Epi::kSlots > 0) p = *(const float4*)(cb + (0 * kC + c) * 512);
                if (Epi::kSlots > 1) a0 = *(const float4*)(cb + (1 * kC + c) * 512);
                if (Epi::kSlots > 2) a1 = *(const float4*)(cb + (2 * kC + c) * 512);
                float4 g = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                                       __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
                epi.apply(crow, D4, gl + c * kG, g, acc[c], p, a0, a1);
              }
            }
            if (gl == 0) epi.touch(crow);
          }
        }
      }
    }
    st_c = st_c + 1 == kNS ? 0 : st_c + 1;
    if (anyv) empties = 0;
    else if (++empties == kNS - 1) break;   // every issued slot has been consumed
  }
}

// shared-memory opt-in of one instantiation, once per device
template <typename Kernel>
int stream_attr_once(Kernel kernel, int smem_bytes, std::atomic<unsigned long long>* seen) {
  int dev = 0;
  EDL_CUDA(cudaGetDevice(&dev));
  EDL_REQUIRE(dev >= 0 && dev < 64, EDL_ELIMIT, "segreduce: device index %d out of range", dev);
  const unsigned long long bit = 1ull << dev;
  if (!(seen->load(std::memory_order_acquire) & bit)) {
    EDL_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    seen->fetch_or(bit, std::memory_order_release);
  }
  return EDL_OK;
}

template <int kG, int kC, typename GT, typename Epi>
int launch_stream_main(const void* grad, int D4, const int32_t* order, const int32_t* seg, const int32_t* seg_first,
                       const int32_t* n_unique, int64_t max_unique, const int64_t* uniq, const int64_t* layout,
                       Epi epi, int32_t* ticket_counter, int long_thr, cudaStream_t st) {
  typedef StreamShape<kG, kC, GT, Epi> S;
  static std::atomic<unsigned long long> seen{0};
  int rc = stream_attr_once(k_seg_stream<kG, kC, GT, Epi>, S::kSmem, &seen);
  if (rc) return rc;
  const int fit = (227 * 1024) / (S::kSmem + 1024);
  const int per_sm = std::max(1, std::min(edl_tune(EDL_TUNE_STREAM_CTAS), fit));
  const int64_t want = edl_cdiv(edl_cdiv(max_unique, 32), S::kWarps);
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)EDL_SM_COUNT * per_sm));
  k_seg_stream<kG, kC, GT, Epi><<<grid, kThreads, S::kSmem, st>>>((const unsigned char*)grad, D4, order, seg, seg_first,
                                                                  n_unique, uniq, layout, epi, ticket_counter, long_thr);
  return EDL_OK;
}

}  // namespace
