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
#include <type_traits>

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
  // ring depth: the deepest of 4 / 3 / 2 stages that still lets two CTAs share an SM (110 KB each)
  static constexpr int kBudget = 108 * 1024;
  static constexpr int kNS = kWarps * 4 * kStageBytes <= kBudget ? 4 : (kWarps * 3 * kStageBytes <= kBudget ? 3 : 2);
  static constexpr int kSmem = kWarps * kNS * kStageBytes + kWarps * 512;   // rings, then 32 x 16-byte id records per warp
  static constexpr int kMinCtas = kC == 1 ? 3 : (kC == 2 ? 2 : 1);
};

// shared-memory accesses by 32-bit shared-window address (no generic-pointer arithmetic in the loop)
__device__ __forceinline__ void cp_async16_s(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ double2 lds_d2(uint32_t a) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long lds_u64(uint32_t a) {
  unsigned long long v;
  asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts_u64(uint32_t a, unsigned long long v) {
  asm volatile("st.shared.u64 [%0], %1;" ::"r"(a), "l"(v) : "memory");
}

template <int kG, int kC, typename GT, typename Epi>
__global__ void __launch_bounds__(kThreads, (StreamShape<kG, kC, GT, Epi>::kMinCtas))
k_seg_stream(const unsigned char* __restrict__ grad, const unsigned char* __restrict__ grad32, int D4,
             const int32_t* __restrict__ order,
             const int32_t* __restrict__ seg, const int32_t* __restrict__ seg_first,
             const int32_t* __restrict__ n_unique, const int64_t* __restrict__ uniq,
             const int64_t* __restrict__ layout, Epi epi, int32_t* __restrict__ ticket_counter, int kLongThr,
             int stages) {   // stages: ring slots in use, 2..kNS (more bytes in flight = longer queues for every dependent access)
  typedef StreamShape<kG, kC, GT, Epi> S;
  typedef Row<GT> R;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  epi.prepare();
  __syncthreads();
  constexpr unsigned kFull = 0xffffffffu;
  constexpr bool kIsMixed = std::is_same<GT, MixedRow>::value;
  constexpr int kNS = S::kNS, kNP = S::kNP, kSlotP = S::kSlotP, kH = S::kH;
  constexpr uint32_t kStage = S::kStageBytes;
  const int U = *n_unique;
  const unsigned lane = threadIdx.x & 31u;
  const int warp = (int)threadIdx.x >> 5;
  const int gl = (int)lane % kG, grp = (int)lane / kG;
  // this lane's 16-byte piece q of stage t lives at pbase + t * kStage + q * 512, its descriptor at dbase + t * kStage
  const uint32_t pbase = smem_u32(smem_raw) + (uint32_t)warp * (kNS * kStage) + lane * 16u;
  const uint32_t dbase = smem_u32(smem_raw) + (uint32_t)warp * (kNS * kStage) + kNP * 512u + lane * 8u;
  const int W = (int)gridDim.x * S::kWarps;
  const size_t ROWB_G = (size_t)D4 * R::kBytes4;   // bytes of a gradient row
  const size_t ROWB = epi.row_bytes();             // bytes from a parameter / slot row to the next
  bool colok[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) colok[c] = gl + c * kG < D4;

  // ---- id records.  A ticket is 32 consecutive unique ids, one per lane; when it becomes current its
  // eligible ids (1..kLongThr gradient rows) are written, compacted, to the warp's record table in shared
  // memory (16 bytes each: first position, first gradient row, row | count << 40), and lane groups that
  // finished an id take the next records in order -- whichever group is free, so a many-row id does not
  // hold up the ids behind it.  The next ticket's raw loads were issued one ticket ago; the ticket after
  // that is still on its way back from the counter.
  const uint32_t rbase = smem_u32(smem_raw) + (uint32_t)S::kWarps * (kNS * kStage) + (uint32_t)warp * 512u;
  int n_elig = 0, next_rec = 0;
  // raw records (segment bounds, first row, token) of the next two tickets, and the two tickets after them
  int a_s0 = 0, a_s1 = 0, a_first = -1, b_s0 = 0, b_s1 = 0, b_first = -1;
  long long a_tok = 0, b_tok = 0;
  bool done = false;         // the ticket just promoted starts beyond U: no work left
  const int gw = (int)blockIdx.x * S::kWarps + warp;
  int tA = gw, tB = gw + W;            // tickets of a_*, b_*: the first four per warp are fixed, no atomic
  int pA = gw + 2 * W, pB = gw + 3 * W;   // lane 0: tickets to load next (pB: the counter's latest answer)
  const unsigned lt_mask = edl_lanemask_lt();

  auto load_raw = [&](int tk, int& s0, int& s1, int& first, long long& tok) {
    const long long u = (long long)tk * 32 + lane;
    s0 = s1 = 0;
    first = -1;
    if (u < U) {
      s0 = __ldg(seg + u);
      s1 = __ldg(seg + u + 1);
      if (seg_first) first = __ldg(seg_first + u);
      tok = epi.token(uniq, (int)u);
    }
  };
  // Ticket a becomes current; b moves up; the ticket the counter handed out two promotions ago is loaded
  // into b; a new ticket is requested.  Everything consumed here was requested at least one whole ticket
  // (8+ iterations) earlier: under load a dependent access takes microseconds, a ticket lasts longer.
  auto promote = [&]() {
    const int cnt = a_s1 - a_s0;
    long long rw = 0;
    bool mine = cnt > 0 && cnt <= kLongThr;
    if (mine) {
      rw = epi.row_from(&L, a_tok);
      mine = rw >= 0;              // a write epilogue may drop a row (beyond its slot)
    }
    int fr = a_first;
    if (mine && fr < 0) fr = __ldg(order + a_s0);
    const unsigned avail = __ballot_sync(kFull, mine);
    n_elig = __popc(avail);
    next_rec = 0;
    if (mine) {
      const uint32_t at = rbase + (uint32_t)__popc(avail & lt_mask) * 16u;
      const unsigned long long packed = (unsigned long long)rw | ((unsigned long long)cnt << 40);
      asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(at), "r"(a_s0), "r"(fr),
                   "r"((unsigned)packed), "r"((unsigned)(packed >> 32)) : "memory");
    }
    __syncwarp();
    done = (long long)tA * 32 >= U;
    a_s0 = b_s0; a_s1 = b_s1; a_first = b_first; a_tok = b_tok; tA = tB;
    tB = __shfl_sync(kFull, pA, 0);
    load_raw(tB, b_s0, b_s1, b_first, b_tok);
    pA = pB;
    if (lane == 0) pB = atomicAdd(ticket_counter, 1) + 4 * W;
  };
  load_raw(tA, a_s0, a_s1, a_first, a_tok);
  load_raw(tB, b_s0, b_s1, b_first, b_tok);
  promote();

  // ---- the lane group's current id
  int pos = 0, end = 0;      // next position in order[] and segment end; pos == end: none
  int j = 0;                 // gradient row of position pos
  long long row = 0;         // epilogue row (slab row / output position), < 2^40
  bool fresh = false;        // the next slot is the id's first
  double acc[kC][4];
#pragma unroll
  for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;

  uint32_t so_i = 0, so_c = 0;   // byte offsets of the stage to issue into / to consume
  stages = stages < 2 ? 2 : (stages > (kNS < 4 ? kNS : 4) ? (kNS < 4 ? kNS : 4) : stages);
  const uint32_t ring_bytes = (uint32_t)stages * kStage;
  int empties = 0;
  for (int it = 0;; ++it) {
    // ---- groups that finished their id take the next records of the ticket, in group order
    const unsigned wmask = __ballot_sync(kFull, pos == end);
    if (wmask) {
      if (next_rec >= n_elig) {                     // every record of the ticket has been handed out
        __syncwarp();                               // ... and read: the table may be rewritten
        while (!done) {
          promote();
          if (n_elig) break;
        }
      }
      const int mine_idx = next_rec + __popc(wmask & ((1u << (grp * kG)) - 1u)) / kG;
      next_rec += __popc(wmask) / kG;
      if (pos == end && mine_idx < n_elig) {
        unsigned r0, r1, r2, r3;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                     : "r"(rbase + (uint32_t)mine_idx * 16u) : "memory");
        pos = (int)r0;
        end = (int)r0 + (int)(r3 >> 8);
        j = (int)r1;
        row = (long long)(((unsigned long long)(r3 & 0xffu) << 32) | r2);
        fresh = true;
      }
    }
    // ---- issue one slot: the gradient row at `pos`; with an id's last row also its parameter / slot rows
    unsigned long long d = 0;
    if (pos < end) {
      const bool last = pos + 1 == end;
      bool narrow = false;   // mixed rows: this one is an fp32 row of grad32 (one piece per column)
      if constexpr (kIsMixed) narrow = (j & kMixed32) != 0;
      if (narrow) {
        const unsigned char* gsrc = grad32 + (size_t)(j & (kMixed32 - 1)) * ((size_t)D4 * 16) + (size_t)gl * 16;
#pragma unroll
        for (int c = 0; c < kC; ++c)
          if (colok[c]) cp_async16_s(pbase + so_i + (kSlotP + c * kH) * 512u, gsrc + (size_t)c * kG * 16);
      } else {
        const unsigned char* gsrc = grad + (size_t)j * ROWB_G + (size_t)gl * R::kBytes4;
#pragma unroll
        for (int c = 0; c < kC; ++c) {
          if (colok[c]) {
#pragma unroll
            for (int h = 0; h < kH; ++h)
              cp_async16_s(pbase + so_i + (kSlotP + c * kH + h) * 512u, gsrc + (size_t)c * kG * R::kBytes4 + h * 16);
          }
        }
      }
      if (Epi::kSlots > 0 && last) {
        const size_t roff = (size_t)row * ROWB + (size_t)gl * 16;
#pragma unroll
        for (int s = 0; s < Epi::kSlots; ++s) {
          const unsigned char* b = (const unsigned char*)epi.base(s) + roff;
#pragma unroll
          for (int c = 0; c < kC; ++c)
            if (colok[c]) cp_async16_s(pbase + so_i + (s * kC + c) * 512u, b + (size_t)c * kG * 16);
        }
      }
      d = 1ull | (fresh ? 2ull : 0ull) | (last ? 4ull : 0ull) | (narrow ? 8ull : 0ull) | ((unsigned long long)row << 8);   // row < 2^40
      fresh = false;
      ++pos;
      if (pos < end) j = __ldg(order + pos);   // needed one iteration from now
    }
    sts_u64(dbase + so_i, d);
    cp_async_commit();
    const unsigned anyv = __ballot_sync(kFull, (d & 1ull) != 0);
    so_i = so_i + kStage == ring_bytes ? 0u : so_i + kStage;
    if (it < stages - 1) continue;   // ring not yet primed
    if (stages >= 4) cp_async_wait<3>();
    else if (stages == 3) cp_async_wait<2>();
    else cp_async_wait<1>();
    // ---- consume the oldest slot
    const unsigned long long cd = lds_u64(dbase + so_c);
    if (cd & 1ull) {
      const uint32_t cb = pbase + so_c;
      const bool first = (cd & 2ull) != 0, last = (cd & 4ull) != 0;
      const long long crow = (long long)(cd >> 8);
      typename R::V v[kC];
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        if (colok[c]) {
          if constexpr (kH == 1) {
            v[c] = lds_f4(cb + (kSlotP + c) * 512u);
          } else {
            if (kIsMixed && (cd & 8ull)) {
              if constexpr (kIsMixed) v[c] = R::widen(lds_f4(cb + (kSlotP + c * 2) * 512u));
            } else {
              v[c].a = lds_d2(cb + (kSlotP + c * 2) * 512u);
              v[c].b = lds_d2(cb + (kSlotP + c * 2 + 1) * 512u);
            }
          }
        }
      }
      if (!(first && last)) {   // more than one gradient row: float64 accumulation in row order
        if (first) {
#pragma unroll
          for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
        }
#pragma unroll
        for (int c = 0; c < kC; ++c)
          if (colok[c]) R::add(acc[c], v[c]);
      }
      if (last) {
#pragma unroll
        for (int c = 0; c < kC; ++c) {
          if (colok[c]) {
            float4 p = make_float4(0, 0, 0, 0), a0 = p, a1 = p;
            if (Epi::kSlots > 0) p = lds_f4(cb + (0 * kC + c) * 512u);
            if (Epi::kSlots > 1) a0 = lds_f4(cb + (1 * kC + c) * 512u);
            if (Epi::kSlots > 2) a1 = lds_f4(cb + (2 * kC + c) * 512u);
            // one epilogue for both kinds of id (a warp usually holds both): an id that occurs once takes
            // its gradient row as the sum, the others the rounded float64 accumulator
            float4 g = R::f32(v[c]);
            if (Epi::kOutF64) {
              if (first) { acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0; R::add(acc[c], v[c]); }
            }
            if (!first)
              g = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                              __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
            if (Epi::kMixed && first) epi.apply_single(crow, D4, gl + c * kG, g);   // one gradient row: sent as it is
            else epi.apply(crow, D4, gl + c * kG, g, acc[c], p, a0, a1);
          }
        }
        if (gl == 0) epi.touch(crow);
      }
    }
    so_c = so_c + kStage == ring_bytes ? 0u : so_c + kStage;
    if (anyv) empties = 0;
    else if (++empties == stages - 1) break;   // every issued slot has been consumed
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

// The register path gives a row one float4 per lane (kC == 1).  The streamed kernel runs the same rows
// with lane groups half as wide and two float4 per lane: twice the ids per warp iteration, so the
// per-iteration bookkeeping (hand-out shuffles, descriptors, address arithmetic) is paid half as often.
template <int kG0, int kC0, typename GT, typename Epi>
int launch_stream_main(const void* grad, const void* grad32, int D4, const int32_t* order, const int32_t* seg, const int32_t* seg_first,
                       const int32_t* n_unique, int64_t max_unique, const int64_t* uniq, const int64_t* layout,
                       Epi epi, int32_t* ticket_counter, int long_thr, cudaStream_t st) {
  constexpr int kG = (kC0 == 1 && kG0 >= 8) ? kG0 / 2 : kG0;
  constexpr int kC = (kC0 == 1 && kG0 >= 8) ? 2 : kC0;
  typedef StreamShape<kG, kC, GT, Epi> S;
  static std::atomic<unsigned long long> seen{0};
  int rc = stream_attr_once(k_seg_stream<kG, kC, GT, Epi>, S::kSmem, &seen);
  if (rc) return rc;
  const int fit = (227 * 1024) / (S::kSmem + 1024);
  const int per_sm = std::max(1, std::min(edl_tune(EDL_TUNE_STREAM_CTAS), fit));
  const int64_t want = edl_cdiv(edl_cdiv(max_unique, 32), S::kWarps);
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)EDL_SM_COUNT * per_sm));
  k_seg_stream<kG, kC, GT, Epi><<<grid, kThreads, S::kSmem, st>>>((const unsigned char*)grad, (const unsigned char*)grad32, D4, order, seg, seg_first,
                                                                  n_unique, uniq, layout, epi, ticket_counter, long_thr,
                                                                  edl_tune(EDL_TUNE_STREAM_STAGES));
  return EDL_OK;
}

}  // namespace
