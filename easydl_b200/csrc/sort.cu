// A1 dedup/unique and A2 hash-mod partition: LSD radix sort built from one stable
// counting-sort pass (per-tile histogram -> per-digit row scan -> ranked scatter), then a
// head-flag scan that emits uniq / inverse / counts / seg_offsets.
//
// SPEC: A1 == np.unique(keys, return_inverse, return_counts) on the fused
// key; A2 == stable partition by (per-table id mod S).  Integer work, bit-exact.
//
// Stability inside a tile comes from warp-striped ranking with __match_any_sync: warp w owns
// 256 consecutive elements, iteration k ranks elements w*256 + k*32 + lane, so rank order ==
// input order.  Everything for a cfg2 batch (1.7M keys, 28 bits) stays L2-resident.
#include <cooperative_groups.h>

#include <algorithm>
#include <mutex>
#include <type_traits>

#include "edl_common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int kThreads = 256;
constexpr int kItems = 8;
constexpr int kTile = kThreads * kItems;  // 2048
constexpr int kWarps = kThreads / 32;

// ------------------------------------------------------------------ loaders / digit functors
template <typename KeyT>
struct IdsLoader {  // pass 0 of the sort: fused key formed on the fly, payload = position
  const int64_t* ids;
  const int64_t* field_base;
  int n_fields;
  __device__ __forceinline__ KeyT key(int, int64_t j) const {
    long long k = ids[j];
    if (field_base) k += field_base[(int)(j % n_fields)];
    return (KeyT)k;
  }
  __device__ __forceinline__ uint32_t val(int, int64_t j) const { return (uint32_t)j; }
  __device__ __forceinline__ int64_t flat(int, int64_t j) const { return j; }
};

// Field-major pass 0 (fused [B, F] id batches): element j' = f * stride + b reads ids[b * F + f].
// Keys are the per-field ids (the field is the segment, see Seg), payload = batch position.
template <typename KeyT>
struct FieldMajorLoader {
  const int64_t* ids;
  int n_fields;
  int64_t stride;   // padded samples per field
  __device__ __forceinline__ KeyT key(int f, int64_t b) const { return (KeyT)ids[b * n_fields + f]; }
  __device__ __forceinline__ uint32_t val(int f, int64_t b) const { return (uint32_t)(b * n_fields + f); }
  __device__ __forceinline__ int64_t flat(int f, int64_t b) const { return (int64_t)f * stride + b; }
};

// Segmented counting sort: the tiles are grouped into independent segments (one per field) that
// are sorted separately and stay in segment order.  tps == 0: one segment, the flat case.
struct Seg {
  int tps;             // tiles per segment (0 = all tiles form one segment)
  int64_t in_stride;   // distance between segment starts in the input index space
  int64_t valid;       // valid elements at the start of every segment
  int64_t out_stride;  // distance between segment starts in the output index space
};

template <typename KeyT>
struct PairLoader {  // later passes
  const KeyT* keys;
  const uint32_t* vals;
  int64_t stride;   // distance between segment starts (0: flat)
  __device__ __forceinline__ KeyT key(int f, int64_t j) const { return keys[(int64_t)f * stride + j]; }
  __device__ __forceinline__ uint32_t val(int f, int64_t j) const { return vals[(int64_t)f * stride + j]; }
  __device__ __forceinline__ int64_t flat(int f, int64_t j) const { return (int64_t)f * stride + j; }
};

struct KeyIndexLoader {  // partition: int64 key, payload = position
  const int64_t* keys;
  __device__ __forceinline__ int64_t key(int, int64_t j) const { return keys[j]; }
  __device__ __forceinline__ uint32_t val(int, int64_t j) const { return (uint32_t)j; }
  __device__ __forceinline__ int64_t flat(int, int64_t j) const { return j; }
};

template <typename KeyT>
struct BitDigit {
  int shift;
  __device__ __forceinline__ unsigned operator()(KeyT k, int64_t) const {
    return (unsigned)(k >> shift) & 0xFFu;
  }
};

struct ArrayDigit {  // partition: digit precomputed by k_shard_digits
  const uint8_t* dig;
  __device__ __forceinline__ unsigned operator()(int64_t, int64_t j) const { return dig[j]; }
};

__device__ __forceinline__ int64_t eff_n(int64_t n, const int32_t* n_dev) {
  if (n_dev) {
    int64_t m = *n_dev;
    return m < n ? m : n;
  }
  return n;
}

// ------------------------------------------------------------------ pass kernel 1: histogram
template <typename KeyT, typename Loader, typename DigitFn>
__global__ void __launch_bounds__(kThreads)
k_hist(Loader ld, DigitFn dg, int64_t n, const int32_t* n_dev, int n_tiles, uint32_t* __restrict__ hist,
       Seg sg) {
  __shared__ unsigned h[256];
  const int tid = threadIdx.x;
  h[tid] = 0;
  __syncthreads();
  const int f = sg.tps ? (int)blockIdx.x / sg.tps : 0;
  const int lt = sg.tps ? (int)blockIdx.x % sg.tps : (int)blockIdx.x;
  const int64_t ne = sg.tps ? sg.valid : eff_n(n, n_dev);
  const int64_t base = (int64_t)lt * kTile;
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t j = base + k * kThreads + tid;
    if (j < ne) atomicAdd(&h[dg(ld.key(f, j), ld.flat(f, j))], 1u);
  }
  __syncthreads();
  hist[(size_t)tid * n_tiles + blockIdx.x] = h[tid];
}

// ------------------------------------------------------------------ pass kernel 2: row scan
// One CTA per digit: exclusive scan of its row of per-tile counts, in place; row total out.
__global__ void __launch_bounds__(kThreads)
k_rowscan(uint32_t* __restrict__ hist, int n_tiles, uint32_t* __restrict__ tot, int tps) {
  __shared__ unsigned warp_sum[kWarps];
  // blockIdx.x = f * 256 + digit; the row piece of segment f is [f * tps, (f + 1) * tps)
  const int d = (int)blockIdx.x & 255, f = (int)blockIdx.x >> 8;
  const int len = tps ? tps : n_tiles;
  uint32_t* row = hist + (size_t)d * n_tiles + (size_t)f * len;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int per = (len + kThreads - 1) / kThreads;
  const int b = tid * per;
  const int e = min(b + per, len);
  unsigned s = 0;
  for (int i = b; i < e; ++i) s += row[i];
  unsigned inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) warp_sum[w] = inc;
  __syncthreads();
  unsigned wbase = 0, total = 0;
#pragma unroll
  for (int i = 0; i < kWarps; ++i) {
    unsigned v = warp_sum[i];
    if (i < w) wbase += v;
    total += v;
  }
  unsigned run = wbase + inc - s;
  for (int i = b; i < e; ++i) {
    unsigned v = row[i];
    row[i] = run;
    run += v;
  }
  if (tid == 0) tot[blockIdx.x] = total;
}

// Segmented variant for short row pieces (tps <= 32: one lane per tile): one WARP per
// (segment, digit) instead of one CTA.  blockDim = 256 -> 8 row pieces per CTA.
__global__ void __launch_bounds__(kThreads)
k_rowscan_warp(uint32_t* __restrict__ hist, int n_tiles, uint32_t* __restrict__ tot, int tps, int n_rows) {
  const int r = (int)blockIdx.x * kWarps + (threadIdx.x >> 5);   // r = f * 256 + digit
  if (r >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int d = r & 255, f = r >> 8;
  uint32_t* row = hist + (size_t)d * n_tiles + (size_t)f * tps;
  const unsigned v = lane < tps ? row[lane] : 0u;
  unsigned inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += x;
  }
  if (lane < tps) row[lane] = inc - v;
  if (lane == 31) tot[r] = inc;
}

// ------------------------------------------------------------------ pass kernel 3: scatter
// All global inputs of the tile (keys, payloads, digit totals, this tile's scanned counts) are
// requested up front so their latencies overlap; digit and rank share one register per key.
// 6 CTAs per SM keep the 832 tiles of a cfg2 batch in a single wave.
template <typename KeyT, typename Loader, typename DigitFn, typename KeyOutT>
__global__ void __launch_bounds__(kThreads, 6)
k_scatter(Loader ld, DigitFn dg, int64_t n, const int32_t* n_dev, int n_tiles,
          const uint32_t* __restrict__ hist, const uint32_t* __restrict__ tot,
          KeyOutT* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
          int32_t* __restrict__ pos_out, int32_t* __restrict__ offsets_out, int n_bins, Seg sg) {
  __shared__ unsigned cnt[kWarps][256];
  __shared__ unsigned warp_sum[kWarps];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int f = sg.tps ? (int)blockIdx.x / sg.tps : 0;
  const int ltile = sg.tps ? (int)blockIdx.x % sg.tps : (int)blockIdx.x;
  const int64_t ne = sg.tps ? sg.valid : eff_n(n, n_dev);
  const int64_t wbeg = (int64_t)ltile * kTile + w * (32 * kItems);
  // ---- issue every global load first
  const unsigned t = tot[f * 256 + tid];
  const unsigned tile_off = hist[(size_t)tid * n_tiles + blockIdx.x];
  KeyT key[kItems];
  uint32_t val[kItems];
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    const int64_t j = wbeg + k * 32 + lane;
    if (j < ne) { key[k] = ld.key(f, j); val[k] = ld.val(f, j); }
    else { key[k] = (KeyT)0; val[k] = 0u; }
  }
#pragma unroll
  for (int i = 0; i < kWarps; ++i) cnt[i][tid] = 0;

  // global digit base: exclusive scan over the 256 digit totals (thread d <-> digit d)
  unsigned inc = t;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) warp_sum[w] = inc;
  __syncthreads();
  unsigned wbase = 0;
#pragma unroll
  for (int i = 0; i < kWarps; ++i)
    if (i < w) wbase += warp_sum[i];
  const unsigned dbase = wbase + inc - t;  // first output slot of digit `tid`
  if (offsets_out && blockIdx.x == 0) {     // partition: offsets[S+1]
    if (tid < n_bins) offsets_out[tid] = (int32_t)dbase;
    if (tid == n_bins) offsets_out[tid] = (int32_t)ne;
  }

  unsigned dr[kItems];   // digit in bits 0..8 (256 = invalid), rank within the warp's digit run above
  const unsigned lt = edl_lanemask_lt();
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    const int64_t j = wbeg + k * 32 + lane;
    const bool valid = j < ne;
    const unsigned d = valid ? dg(key[k], ld.flat(f, j)) : 256u;
    const unsigned mask = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(mask) - 1;
    unsigned prev = 0;
    if (lane == leader && valid) prev = cnt[w][d];
    prev = __shfl_sync(0xffffffffu, prev, leader);
    dr[k] = d | ((prev + __popc(mask & lt)) << 9);
    if (lane == leader && valid) cnt[w][d] = prev + __popc(mask);
    __syncwarp();
  }
  __syncthreads();
  {
    unsigned run = dbase + tile_off + (unsigned)((int64_t)f * sg.out_stride);
#pragma unroll
    for (int i = 0; i < kWarps; ++i) {
      unsigned c = cnt[i][tid];
      cnt[i][tid] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    const unsigned d = dr[k] & 511u;
    if (d < 256u) {
      const unsigned p = cnt[w][d] + (dr[k] >> 9);
      keys_out[p] = (KeyOutT)key[k];
      vals_out[p] = val[k];
      if (pos_out) pos_out[val[k]] = (int32_t)p;
    }
  }
}

// ------------------------------------------------------------------ field sort inside one cluster
// Field-major batches whose fields hold at most 65536 samples: the whole LSD sort of ONE field runs
// inside ONE thread-block cluster, the field's (key, sample) pairs never leaving shared memory between
// the passes.  CTA r of the cluster holds positions [r * chunk, (r + 1) * chunk) of the field; a pass ranks
// its elements like k_scatter does (warp-striped, stable), groups them by digit in a second local buffer,
// publishes its 256-bin histogram, reads the other CTAs' histograms through distributed shared memory, and
// copies every digit group into the shared memory of the CTA(s) that own its destination.  One launch replaces
// the 3 x (histogram, row scan, scatter) launches of the global path and, in its fused form, the three kernels
// of the unique step as well (otherwise the head count, when the tiles line up).
// Two shapes (template parameters kCS = CTAs per cluster, kCT = threads per CTA; knob dedup_shape):
//   4 x 1024: 16384 positions and 215 KB of shared memory per CTA, one CTA per SM, the whole register file
//   8 x 512 :  8192 positions and 107 KB per CTA, half the register file: other kernels' CTAs fit beside it
constexpr int kCIter = 16;                   // 32-element rounds per warp and pass, at most
constexpr int kClusterMaxField = 65536;      // kCS * (kCT / 32) * 32 * kCIter for both shapes

static inline size_t cluster_sort_smem(int chunk, int warps) {
  return (size_t)chunk * 12 + (size_t)warps * 256 * 2 + 256 * 4 * 3 + 64 + 64 + ((size_t)chunk * 4 >= (size_t)warps * 2048 ? 0 : (size_t)warps * 2048);
}

// What the fused form of k_cluster_sort writes besides `order` (the whole unique step).
struct FusedUnique {
  int on;
  int64_t* uniq; int32_t* seg_offsets; int32_t* seg_first; int32_t* inverse; int32_t* n_unique;
  const int64_t* field_base;
  unsigned* ticket;                 // zeroed before the launch: fields are handed out in arrival order
  unsigned long long* field_heads;  // [F], zeroed: (1 << 32) | distinct ids of the field, once known
};

// ids [B, F] int64 -> keys [F, B] uint32, both sides in whole lines through a shared-memory tile.  Read straight
// from the id matrix, the ids of one field are 8 bytes out of every F * 8: one sector per id, and the sort
// kernel's load phase was bound by outstanding sectors (ncu: a quarter of its samples).
constexpr int kTS = 128;            // samples per CTA
__global__ void __launch_bounds__(256)
k_field_major_keys(const int64_t* __restrict__ ids, int F, int B, uint32_t* __restrict__ keys) {
  extern __shared__ uint32_t tile[];   // [F][kTS + 1]
  const int b0 = (int)blockIdx.x * kTS;
  const int nb = min(kTS, B - b0);
  const int64_t* src = ids + (int64_t)b0 * F;
  const int total = nb * F;
  for (int i = threadIdx.x; i < total; i += 256) {
    const int b = i / F, f = i - b * F;
    tile[f * (kTS + 1) + b] = (uint32_t)src[i];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < F * kTS; i += 256) {
    const int f = i / kTS, b = i % kTS;
    if (b < nb) keys[(int64_t)f * B + b0 + b] = tile[f * (kTS + 1) + b];
  }
}

template <int kCS, int kCT>
__global__ void __cluster_dims__(kCS, 1, 1) __launch_bounds__(kCT, 1024 / kCT)   // 64 registers per thread either way
k_cluster_sort(const int64_t* __restrict__ ids, const uint32_t* __restrict__ keys_fm, int F, int B, int chunk,
               int passes, uint32_t* __restrict__ keys_out, int32_t* __restrict__ order,
               uint32_t* __restrict__ tile_heads, FusedUnique fu) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ unsigned s_misc[4];    // [0] field of this cluster (rank 0's copy), [1] heads of this CTA, [2] ids before this field
  constexpr int kCW = kCT / 32;
  __shared__ unsigned s_wsum[32];
  cg::cluster_group cluster = cg::this_cluster();
  const int r = (int)cluster.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  int f = (int)blockIdx.x / kCS;
  if (fu.on) {
    // Fields in ARRIVAL order: a cluster waits (below) only for clusters that took their ticket before it did,
    // and those are running.
    if (r == 0 && tid == 0) s_misc[0] = atomicAdd(fu.ticket, 1u);
    cluster.sync();
    f = (int)*cluster.map_shared_rank(&s_misc[0], 0);
  }
  uint32_t* kbuf0 = (uint32_t*)smem_raw;
  uint32_t* kbuf1 = kbuf0 + chunk;
  uint16_t* vbuf0 = (uint16_t*)(kbuf1 + chunk);
  uint16_t* vbuf1 = vbuf0 + chunk;
  uint16_t* cnt = vbuf1 + chunk;                       // [kCW][256]
  uint32_t* hist = (uint32_t*)(cnt + kCW * 256);       // [256], read by the other CTAs
  uint32_t* base = hist + 256;                         // [256]
  uint32_t* lstart = base + 256;                       // [256]
  uint32_t* warp_sum = lstart + 256;                   // [8] (+ pad)
  uint32_t* tile_cnt = warp_sum + 16;                  // [8] (+ pad)
  // [kCW][2][256] peer-mask words of the ranking: in kbuf1 when that is large enough, else behind everything
  uint32_t* mm_base = chunk * 4 >= kCW * 2048 ? kbuf1 : tile_cnt + 16;
  const int iters = chunk / (kCW * 32);
  const int p0 = r * chunk;                            // first field position this CTA holds

  {                        // the field's ids of this CTA's positions: every load of a thread in flight at once
    long long raw[kCIter];
#pragma unroll
    for (int k = 0; k < kCIter; ++k) {
      const int p = p0 + k * kCT + tid;
      raw[k] = 0ll;
      if (k < iters && p < B) raw[k] = keys_fm ? (long long)keys_fm[(int64_t)f * B + p] : ids[(int64_t)p * F + f];
    }
#pragma unroll
    for (int k = 0; k < kCIter; ++k) {
      const int e = k * kCT + tid;
      if (k < iters && p0 + e < B) {
        kbuf0[e] = (uint32_t)raw[k];
        vbuf0[e] = (uint16_t)(p0 + e);
      }
    }
  }
  if (tid < 8) tile_cnt[tid] = 0;
  __syncthreads();

  const unsigned lt = edl_lanemask_lt();
  const int n_local = min(max(B - p0, 0), chunk);      // valid positions of this CTA (a prefix)
  const unsigned kloc = (unsigned)__cvta_generic_to_shared(kbuf0), vloc = (unsigned)__cvta_generic_to_shared(vbuf0);
  for (int pass = 0; pass < passes; ++pass) {
    // Every pass: kbuf0 / vbuf0 -> (ranked, grouped by digit) kbuf1 / vbuf1 of THIS CTA -> kbuf0 / vbuf0 of the
    // CTAs that own the destinations.  Grouping locally first makes the stores into the other CTAs' shared
    // memory run over consecutive addresses (scattered 4-byte remote stores took 3x the time of the rest).
    const int shift = pass * 8;
    uint16_t* wc = cnt + w * 256;
#pragma unroll
    for (int i = 0; i < 8; ++i) wc[i * 32 + lane] = 0;
    __syncwarp();
    // Lanes with my digit: every lane ORs its bit into the digit's word of a per-warp table and reads the word
    // back (8 ballots per round cost 90 instructions, MATCH.ANY walks the warp's distinct values one by one: ncu
    // had either at a third of this kernel).  Two tables in turn, so that the leader's clearing of one round's
    // words cannot meet the next round's ORs.  The tables live in the grouping buffer, idle at this point.
    uint32_t* mm = mm_base + w * 512;
    {
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
      for (int i = 0; i < 4; ++i) ((uint4*)mm)[i * 32 + lane] = z;
    }
    __syncwarp();
    unsigned dr[kCIter];   // digit in bits 0..8 (256 = no element), rank inside the warp's run of the digit above
#pragma unroll
    for (int k = 0; k < kCIter; ++k) {
      dr[k] = 256u;
      if (k < iters) {
        const int e = w * (iters * 32) + k * 32 + lane;
        const bool valid = e < n_local;
        const unsigned d = valid ? ((kbuf0[e] >> shift) & 0xFFu) : 256u;
        uint32_t* m = mm + (k & 1) * 256;
        if (valid) atomicOr(m + d, 1u << lane);
        __syncwarp();
        unsigned peers = 0, prev = 0;
        if (valid) { peers = m[d]; prev = wc[d]; }
        __syncwarp();
        const unsigned below = __popc(peers & lt);
        if (valid && below == 0) {          // first lane of the digit: next round's count, word back to zero
          m[d] = 0u;
          wc[d] = (uint16_t)(prev + __popc(peers));
        }
        dr[k] = d | ((prev + below) << 9);
      }
    }
    __syncthreads();
    unsigned mine = 0;
    if (tid < 256) {       // this CTA's histogram; the warps' counts become their offsets inside the digit
      unsigned run = 0;
#pragma unroll 8
      for (int i = 0; i < kCW; ++i) {
        const unsigned c = cnt[i * 256 + tid];
        cnt[i * 256 + tid] = (uint16_t)run;
        run += c;
      }
      hist[tid] = run;
      mine = run;
    }
    {                      // where each digit starts inside this CTA: exclusive scan of its histogram
      unsigned inc = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
      }
      if (tid < 256 && lane == 31) warp_sum[w] = inc;
      __syncthreads();
      if (tid < 256) {
        unsigned wbase = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (i < w) wbase += warp_sum[i];
        lstart[tid] = wbase + inc - mine;
      }
      __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kCIter; ++k) {
      const unsigned d = dr[k] & 511u;
      if (k < iters && d < 256u) {
        const int e = w * (iters * 32) + k * 32 + lane;
        const unsigned q = lstart[d] + wc[d] + (dr[k] >> 9);
        kbuf1[q] = kbuf0[e];
        vbuf1[q] = vbuf0[e];
      }
    }
    cluster.sync();        // histograms published, every CTA has left its kbuf0 / vbuf0 (and is running)
    unsigned total = 0, below = 0;
    if (tid < 256) {
#pragma unroll
      for (int rr = 0; rr < kCS; ++rr) {
        const uint32_t* h = cluster.map_shared_rank(hist, rr);
        const unsigned v = h[tid];
        total += v;
        if (rr < r) below += v;
      }
    }
    unsigned inc = total;  // exclusive scan of the 256 digit totals (warps 0..7)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (tid < 256 && lane == 31) warp_sum[w] = inc;
    __syncthreads();
    if (tid < 256) {
      unsigned wbase = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (i < w) wbase += warp_sum[i];
      // position inside the field of this CTA's first element with digit `tid`, minus where that element sits here
      base[tid] = wbase + inc - total + below - lstart[tid];
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < kCIter; h += 8) {    // eight elements per thread in flight
      uint32_t ck[8], cp[8];
      uint16_t cv[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int q = (h + k) * kCT + tid;
        if (h + k < iters && q < n_local) {
          ck[k] = kbuf1[q];
          cv[k] = vbuf1[q];
          cp[k] = base[(ck[k] >> shift) & 0xFFu] + (unsigned)q;
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int q = (h + k) * kCT + tid;
        if (h + k < iters && q < n_local) {
          const unsigned p = cp[k];
          unsigned rr = 0;
#pragma unroll
          for (int i = 1; i < kCS; ++i) rr += p >= (unsigned)(i * chunk) ? 1u : 0u;
          const unsigned off = p - rr * chunk;
          unsigned ka, va;       // the same offsets in CTA rr's shared memory (shared::cluster window)
          asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ka) : "r"(kloc + off * 4u), "r"(rr));
          asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(va) : "r"(vloc + off * 2u), "r"(rr));
          asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(ka), "r"(ck[k]) : "memory");
          asm volatile("st.shared::cluster.u16 [%0], %1;" ::"r"(va), "h"(cv[k]) : "memory");
        }
      }
    }
    cluster.sync();        // all stores of the pass have landed; kbuf1 / vbuf1 may be overwritten
  }

  const uint32_t* kf = kbuf0;
  const uint16_t* vf = vbuf0;
  uint32_t prevb = 0;      // the key in front of this CTA's first position
  if (r > 0 && p0 < B) prevb = cluster.map_shared_rank(kf, r - 1)[chunk - 1];

  if (fu.on) {
    // The unique step in place: head flags as one ballot per round, heads counted over warp -> CTA -> cluster ->
    // the fields in front (their totals are published in field_heads as soon as they are known), then uniq /
    // seg_offsets / seg_first / inverse / order are written from shared memory.
    unsigned hb[kCIter];
    unsigned wtot = 0;
#pragma unroll
    for (int k = 0; k < kCIter; ++k) {
      hb[k] = 0u;
      if (k < iters) {
        const int e = w * (iters * 32) + k * 32 + lane;
        bool head = false;
        if (e < n_local) head = (p0 + e == 0) || kf[e] != (e > 0 ? kf[e - 1] : prevb);
        hb[k] = __ballot_sync(0xffffffffu, head);
        wtot += __popc(hb[k]);
      }
    }
    if (lane == 0) s_wsum[w] = wtot;
    __syncthreads();
    if (w == 0) {
      const unsigned v = lane < kCW ? s_wsum[lane] : 0u;
      unsigned inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned x = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += x;
      }
      s_wsum[lane] = inc - v;
      if (lane == 31) s_misc[1] = inc;
    }
    cluster.sync();        // every CTA's head count is known
    unsigned below = 0, field_total = 0;
#pragma unroll
    for (int rr = 0; rr < kCS; ++rr) {
      const unsigned v = *cluster.map_shared_rank(&s_misc[1], rr);
      field_total += v;
      if (rr < r) below += v;
    }
    if (r == 0 && tid == 0)
      atomicExch(fu.field_heads + f, (1ull << 32) | (unsigned long long)field_total);
    if (w == 0) {          // distinct ids of the fields in front of this one
      unsigned sum = 0;
      for (int j = lane; j < f; j += 32) {
        unsigned long long v;
        do { v = *(volatile unsigned long long*)(fu.field_heads + j); } while ((v >> 32) == 0ull);
        sum += (unsigned)v;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      if (lane == 0) s_misc[2] = sum;
    }
    __syncthreads();
    const unsigned before = s_misc[2];
    unsigned run = before + below + s_wsum[w];
    const long long fbase = fu.field_base[f];
    const unsigned le = lt | (1u << lane);
#pragma unroll
    for (int k = 0; k < kCIter; ++k) {
      if (k < iters) {
        const int e = w * (iters * 32) + k * 32 + lane;
        if (e < n_local) {
          const unsigned u = run + __popc(hb[k] & le) - 1u;
          const int64_t i = (int64_t)f * B + p0 + e;
          const int32_t pos = (int32_t)((uint32_t)vf[e] * (uint32_t)F + (uint32_t)f);
          order[i] = pos;
          if (fu.inverse) fu.inverse[pos] = (int32_t)u;
          if ((hb[k] >> lane) & 1u) {
            fu.uniq[u] = (int64_t)kf[e] + fbase;
            fu.seg_offsets[u] = (int32_t)i;
            if (fu.seg_first) fu.seg_first[u] = pos;
          }
        }
        run += __popc(hb[k]);
      }
    }
    if (f == F - 1 && r == 0 && tid == 0) {
      const unsigned U = before + field_total;
      *fu.n_unique = (int32_t)U;
      fu.seg_offsets[U] = (int32_t)((int64_t)F * B);
    }
    cluster.sync();        // no CTA leaves while its neighbours may still read its last key / head count
    return;
  }

  for (int e = tid; e < chunk; e += kCT) {
    const int p = p0 + e;
    unsigned c = 0;
    if (p < B) {
      const uint32_t key = kf[e];
      const uint32_t pv = e > 0 ? kf[e - 1] : prevb;
      const int64_t i = (int64_t)f * B + p;
      keys_out[i] = key;
      order[i] = (int32_t)((uint32_t)vf[e] * (uint32_t)F + (uint32_t)f);
      c = (p == 0 || key != pv) ? 1u : 0u;
    }
    if (tile_heads) {      // chunk and B are multiples of the 2048-element tiles of the unique kernels
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
      if (lane == 0 && c) atomicAdd(&tile_cnt[e >> 11], c);
    }
  }
  __syncthreads();
  if (tile_heads && tid < (chunk >> 11) && p0 + tid * 2048 < B)
    tile_heads[((int64_t)f * B + p0 + tid * 2048) >> 11] = tile_cnt[tid];
  cluster.sync();          // no CTA leaves while its neighbour may still read its last key
}

// ------------------------------------------------------------------ unique kernels
template <typename KeyT>
__global__ void __launch_bounds__(kThreads)
k_head_count(const KeyT* __restrict__ keys, int64_t n, const int32_t* n_dev, uint32_t* __restrict__ tile_heads,
             int64_t seg_len) {
  __shared__ unsigned warp_sum[kWarps];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int64_t ne = eff_n(n, n_dev);
  const int64_t base = (int64_t)blockIdx.x * kTile;
  unsigned c = 0;
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + k * kThreads + tid;
    if (i < ne) c += (i == 0 || keys[i] != keys[i - 1] || (seg_len && i % seg_len == 0)) ? 1u : 0u;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) warp_sum[w] = c;
  __syncthreads();
  if (tid == 0) {
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < kWarps; ++i) s += warp_sum[i];
    tile_heads[blockIdx.x] = s;
  }
}

// single CTA: exclusive scan over the tile head counts; publishes U and seg_offsets[U] = n
__global__ void __launch_bounds__(1024)
k_scan_tiles(uint32_t* __restrict__ tile_heads, int n_tiles, int64_t n, const int32_t* n_dev,
             int32_t* __restrict__ n_unique, int32_t* __restrict__ seg_offsets) {
  __shared__ unsigned warp_sum[32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int per = (n_tiles + 1023) / 1024;
  const int b = tid * per, e = min(b + per, n_tiles);
  unsigned s = 0;
  for (int i = b; i < e; ++i) s += tile_heads[i];
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
    unsigned v = tile_heads[i];
    tile_heads[i] = run;
    run += v;
  }
  if (tid == 0) {
    *n_unique = (int32_t)total;
    seg_offsets[total] = (int32_t)eff_n(n, n_dev);
  }
}

template <typename KeyT>
__global__ void __launch_bounds__(kThreads)
k_unique_write(const KeyT* __restrict__ keys, const uint32_t* __restrict__ order, int64_t n,
               const int32_t* n_dev, const uint32_t* __restrict__ tile_base,
               int64_t* __restrict__ uniq, int32_t* __restrict__ seg_offsets,
               int32_t* __restrict__ seg_first, int32_t* __restrict__ inverse, int64_t seg_len,
               const int64_t* __restrict__ field_base) {
  __shared__ unsigned warp_sum[kWarps];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int64_t ne = eff_n(n, n_dev);
  const int64_t base = (int64_t)blockIdx.x * kTile + (int64_t)tid * kItems;  // blocked
  KeyT key[kItems];
  unsigned flag[kItems];
  unsigned c = 0;
  KeyT prev = 0;
  if (base > 0 && base <= ne) prev = keys[base - 1];
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + k;
    bool valid = i < ne;
    key[k] = valid ? keys[i] : (KeyT)0;
    flag[k] = (valid && (i == 0 || key[k] != prev || (seg_len && i % seg_len == 0))) ? 1u : 0u;
    prev = key[k];
    c += flag[k];
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
#pragma unroll
  for (int i = 0; i < kWarps; ++i)
    if (i < w) wbase += warp_sum[i];
  unsigned run = tile_base[blockIdx.x] + wbase + inc - c;  // heads strictly before my items
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + k;
    if (i < ne) {
      run += flag[k];
      unsigned u = run - 1;
      if (flag[k]) {
        uniq[u] = (int64_t)key[k] + (seg_len ? field_base[i / seg_len] : 0);
        seg_offsets[u] = (int32_t)i;
        if (seg_first) seg_first[u] = (int32_t)order[i];
      }
      if (inverse) inverse[order[i]] = (int32_t)u;
    }
  }
}

__global__ void k_counts(const int32_t* __restrict__ seg_offsets, const int32_t* __restrict__ n_unique,
                         int32_t* __restrict__ counts) {
  const int U = *n_unique;
  for (int u = blockIdx.x * blockDim.x + threadIdx.x; u < U; u += gridDim.x * blockDim.x)
    counts[u] = seg_offsets[u + 1] - seg_offsets[u];
}

// A2: digit = (per-table id) mod S
__global__ void __launch_bounds__(kThreads)
k_shard_digits(const int64_t* __restrict__ keys, int64_t n, const int32_t* n_dev,
               const int64_t* __restrict__ layout, unsigned S, uint8_t* __restrict__ dig) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  const int64_t ne = eff_n(n, n_dev);
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < ne;
       j += (int64_t)gridDim.x * blockDim.x) {
    int t; unsigned long long id;
    edl_split_key(&L, keys[j], &t, &id);
    dig[j] = (uint8_t)(id % S);
  }
}

__global__ void k_empty_unique(int32_t* n_unique, int32_t* seg_offsets) {
  *n_unique = 0;
  seg_offsets[0] = 0;
}

// ------------------------------------------------------------------ workspace carving
struct UniqueWs {
  void* keysA; void* keysB; uint32_t* valsA; uint32_t* valsB; uint32_t* hist; uint32_t* tot; uint32_t* tile_heads;
  size_t bytes;
};

// Field-major mode (fused [B, F] batches): sort every field's ids separately on their own
// `field_bits` instead of the fused key's `key_bits` -- the field is known from the position, so
// its bits need no radix pass (cfg2: 24 instead of 28 bits = 3 passes instead of 4).
struct FmPlan { bool on; int F; int64_t B, Bp; int tpf; bool in_cluster; };

FmPlan fm_plan(int64_t n, const int32_t* n_dev, const int64_t* field_base, int n_fields, int key_bits,
               int field_bits) {
  FmPlan f{false, 1, n, n, 0, false};
  if (field_bits <= 0 || !field_base || n_fields <= 1 || n_dev || n % n_fields != 0) return f;
  // worth it when a radix pass is saved, or when the fields are small enough for k_cluster_sort
  const bool in_cluster = field_bits <= 32 && n / n_fields <= (int64_t)kClusterMaxField && edl_tune(EDL_TUNE_DEDUP_IMPL) >= 1;
  if ((field_bits + 7) / 8 >= (key_bits + 7) / 8 && !in_cluster) return f;
  f.on = true;
  f.in_cluster = in_cluster;
  f.F = n_fields;
  f.B = n / n_fields;
  f.tpf = (int)edl_cdiv(f.B, kTile);
  f.Bp = (int64_t)f.tpf * kTile;
  return f;
}

UniqueWs carve_unique(void* ws, int64_t n, int key_bits, int n_fields, int field_bits) {
  // sized for either mode: field-major pads every field to whole tiles
  int64_t cap = n;
  int64_t tiles = edl_cdiv(n > 0 ? n : 1, kTile);
  int tot_rows = 1;
  bool fm_possible = false;
  if (field_bits > 0 && n_fields > 1 && n % n_fields == 0) {
    fm_possible = true;
    const int64_t tpf = edl_cdiv(n / n_fields, kTile);
    cap = std::max<int64_t>(cap, tpf * kTile * n_fields);
    tiles = std::max<int64_t>(tiles, tpf * n_fields);
    tot_rows = n_fields;
  }
  const size_t kb = (key_bits <= 32 || (field_bits > 0 && field_bits <= 32 && n_fields > 1)) && key_bits <= 32 ? 4 : 8;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = edl_align_up(off + bytes, 256); return o; };
  // the flat sort ping-pongs its payload between `order` and valsB; only the padded field-major
  // sort needs a second scratch payload buffer
  size_t oA = take((size_t)cap * kb), oB = take((size_t)cap * kb), oVA = take(fm_possible ? (size_t)cap * 4 : 0), oVB = take((size_t)cap * 4);
  size_t oH = take((size_t)256 * tiles * 4), oT = take((size_t)256 * 4 * tot_rows), oC = take((size_t)(tiles + 1) * 4);
  UniqueWs r;
  char* p = (char*)ws;
  r.keysA = p + oA; r.keysB = p + oB; r.valsA = (uint32_t*)(p + oVA); r.valsB = (uint32_t*)(p + oVB);
  r.hist = (uint32_t*)(p + oH); r.tot = (uint32_t*)(p + oT); r.tile_heads = (uint32_t*)(p + oC);
  r.bytes = off;
  return r;
}

void rowscan(uint32_t* hist, int n_tiles, uint32_t* tot, int tps, int n_seg, cudaStream_t st) {
  if (tps > 0 && tps <= 32)
    k_rowscan_warp<<<(256 * n_seg + kWarps - 1) / kWarps, kThreads, 0, st>>>(hist, n_tiles, tot, tps, 256 * n_seg);
  else
    k_rowscan<<<256 * n_seg, kThreads, 0, st>>>(hist, n_tiles, tot, tps);
}

// k_cluster_sort needs more dynamic shared memory than the default limit: raised once per device and shape.
template <int kCS, int kCT>
int cluster_sort_prepare() {
  static std::mutex mu;
  static bool done[64] = {false};
  int dev = -1;
  EDL_CUDA(cudaGetDevice(&dev));
  EDL_REQUIRE(dev >= 0 && dev < 64, EDL_ELIMIT, "edl_unique: device ordinal %d beyond 64", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!done[dev]) {
    EDL_CUDA(cudaFuncSetAttribute(k_cluster_sort<kCS, kCT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)cluster_sort_smem(kClusterMaxField / kCS, kCT / 32)));
    done[dev] = true;
  }
  return EDL_OK;
}

template <int kCS, int kCT>
int cluster_sort_launch(const int64_t* ids, const uint32_t* keys_fm, int F, int64_t B, int passes, uint32_t* keys_out,
                        int32_t* order, uint32_t* tile_heads, bool* heads_done, const FusedUnique& fu, cudaStream_t st) {
  const int chunk = (int)(edl_cdiv(edl_cdiv(B, kCS), kCT) * kCT);
  int rc = cluster_sort_prepare<kCS, kCT>();
  if (rc) return rc;
  *heads_done = !fu.on && B % kTile == 0 && chunk % kTile == 0;
  k_cluster_sort<kCS, kCT><<<F * kCS, kCT, cluster_sort_smem(chunk, kCT / 32), st>>>(
      ids, keys_fm, F, (int)B, chunk, passes, keys_out, order, *heads_done ? tile_heads : (uint32_t*)nullptr, fu);
  EDL_LAUNCH_CHECK("k_cluster_sort");
  return EDL_OK;
}

template <typename KeyT>
int unique_impl(const int64_t* ids, int64_t n, const int32_t* n_dev, const int64_t* field_base,
                int n_fields, int key_bits, int field_bits, int64_t* uniq, int32_t* inverse,
                int32_t* counts, int32_t* order, int32_t* seg_offsets, int32_t* seg_first,
                int32_t* n_unique, const UniqueWs& W, cudaStream_t st) {
  const FmPlan fm = fm_plan(n, n_dev, field_base, n_fields, key_bits, field_bits);
  const int sort_tiles = fm.on ? fm.F * fm.tpf : (int)edl_cdiv(n, kTile);
  const int passes = ((fm.on ? field_bits : key_bits) + 7) / 8;
  const int n_seg = fm.on ? fm.F : 1;
  KeyT* kbuf[2] = {(KeyT*)W.keysA, (KeyT*)W.keysB};
  uint32_t* vbuf[2] = {W.valsA, W.valsB};
  if (!fm.on) {   // flat: `order` is one of the two payload buffers, placed so the last pass lands in it
    vbuf[(passes - 1) & 1] = (uint32_t*)order;
    vbuf[passes & 1] = W.valsB;
  }
  const KeyT* sorted = kbuf[(passes - 1) & 1];
  bool heads_done = false;
  bool in_cluster = false;
  if constexpr (std::is_same<KeyT, uint32_t>::value) {
    if (fm.on && fm.in_cluster) {
      const int mode = edl_tune(EDL_TUNE_DEDUP_IMPL);
      // ids -> field-major 32-bit keys first (in the second key buffer, which this path does not use otherwise)
      const uint32_t* keys_fm = nullptr;
      const size_t tile_bytes = (size_t)fm.F * (kTS + 1) * 4;
      if (tile_bytes <= 48 * 1024) {
        k_field_major_keys<<<(unsigned)edl_cdiv(fm.B, kTS), 256, tile_bytes, st>>>(ids, fm.F, (int)fm.B, (uint32_t*)kbuf[1]);
        keys_fm = (const uint32_t*)kbuf[1];
      }
      FusedUnique fu{};
      if (mode == 2) {     // the unique step inside the sort kernel: ticket + per-field totals start at zero
        const size_t sync_bytes = 8 + 8 * (size_t)fm.F;
        EDL_CUDA(cudaMemsetAsync(W.tot, 0, sync_bytes, st));
        fu = FusedUnique{1, uniq, seg_offsets, seg_first, inverse, n_unique, field_base, (unsigned*)W.tot,
                         (unsigned long long*)(W.tot + 2)};
      }
      const int rc = edl_tune(EDL_TUNE_DEDUP_SHAPE) == 1
                         ? cluster_sort_launch<8, 512>(ids, keys_fm, fm.F, fm.B, passes, (uint32_t*)kbuf[0], order,
                                                       W.tile_heads, &heads_done, fu, st)
                         : cluster_sort_launch<4, 1024>(ids, keys_fm, fm.F, fm.B, passes, (uint32_t*)kbuf[0], order,
                                                        W.tile_heads, &heads_done, fu, st);
      if (rc) return rc;
      if (fu.on) {
        if (counts) {
          int g = (int)std::min<int64_t>(edl_cdiv(n, 256), EDL_SM_COUNT * 8);
          k_counts<<<g, 256, 0, st>>>(seg_offsets, n_unique, counts);
          EDL_LAUNCH_CHECK("k_counts");
        }
        return EDL_OK;
      }
      sorted = kbuf[0];
      in_cluster = true;
    }
  }
  for (int p = 0; p < passes && !in_cluster; ++p) {
    const bool last = p == passes - 1;
    BitDigit<KeyT> dg{p * 8};
    KeyT* ko = kbuf[p & 1];
    uint32_t* vo = last ? (uint32_t*)order : vbuf[p & 1];   // the last pass lands the payload in `order`
    Seg sg{0, 0, 0, 0};
    if (fm.on) sg = Seg{fm.tpf, fm.Bp, fm.B, last ? fm.B : fm.Bp};
    if (p == 0 && fm.on) {
      FieldMajorLoader<KeyT> ld{ids, fm.F, fm.Bp};
      k_hist<KeyT><<<sort_tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, sort_tiles, W.hist, sg);
      rowscan(W.hist, sort_tiles, W.tot, sg.tps, n_seg, st);
      k_scatter<KeyT><<<sort_tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, sort_tiles, W.hist, W.tot, ko, vo,
                                                       (int32_t*)nullptr, (int32_t*)nullptr, 0, sg);
    } else if (p == 0) {
      IdsLoader<KeyT> ld{ids, field_base, n_fields};
      k_hist<KeyT><<<sort_tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, sort_tiles, W.hist, sg);
      rowscan(W.hist, sort_tiles, W.tot, sg.tps, n_seg, st);
      k_scatter<KeyT><<<sort_tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, sort_tiles, W.hist, W.tot, ko, vo,
                                                       (int32_t*)nullptr, (int32_t*)nullptr, 0, sg);
    } else {
      PairLoader<KeyT> ld{kbuf[(p - 1) & 1], vbuf[(p - 1) & 1], fm.on ? fm.Bp : 0};
      k_hist<KeyT><<<sort_tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, sort_tiles, W.hist, sg);
      rowscan(W.hist, sort_tiles, W.tot, sg.tps, n_seg, st);
      k_scatter<KeyT><<<sort_tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, sort_tiles, W.hist, W.tot, ko, vo,
                                                       (int32_t*)nullptr, (int32_t*)nullptr, 0, sg);
    }
  }
  EDL_LAUNCH_CHECK("radix passes");
  const int tiles = (int)edl_cdiv(n, kTile);
  const int64_t seg_len = fm.on ? fm.B : 0;
  if (!heads_done) k_head_count<KeyT><<<tiles, kThreads, 0, st>>>(sorted, n, n_dev, W.tile_heads, seg_len);
  k_scan_tiles<<<1, 1024, 0, st>>>(W.tile_heads, tiles, n, n_dev, n_unique, seg_offsets);
  k_unique_write<KeyT><<<tiles, kThreads, 0, st>>>(sorted, (const uint32_t*)order, n, n_dev, W.tile_heads, uniq,
                                                   seg_offsets, seg_first, inverse, seg_len, field_base);
  if (counts) {
    int g = (int)std::min<int64_t>(edl_cdiv(n, 256), EDL_SM_COUNT * 8);
    k_counts<<<g, 256, 0, st>>>(seg_offsets, n_unique, counts);
  }
  EDL_LAUNCH_CHECK("unique kernels");
  return EDL_OK;
}

// key width of the sort: 32-bit keys when the bits that are actually sorted fit
bool sort_keys_32(int64_t n, const int32_t* n_dev, const int64_t* field_base, int n_fields, int key_bits,
                  int field_bits) {
  const FmPlan fm = fm_plan(n, n_dev, field_base, n_fields, key_bits, field_bits);
  return (fm.on ? field_bits : key_bits) <= 32;
}

}  // namespace

extern "C" size_t edl_unique_workspace_bytes(int64_t n, int key_bits, int n_fields, int field_bits) {
  if (n < 0) return 0;
  return carve_unique(nullptr, n, key_bits, n_fields, field_bits).bytes;
}

extern "C" int edl_unique(const int64_t* ids, int64_t n, const int32_t* n_dev,
                          const int64_t* field_base, int n_fields, int key_bits, int field_bits,
                          int64_t* uniq, int32_t* inverse, int32_t* counts, int32_t* order,
                          int32_t* seg_offsets, int32_t* seg_first, int32_t* n_unique,
                          void* workspace, size_t workspace_bytes, edl_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EDL_REQUIRE(n >= 0 && n < (int64_t)0x7fff0000, EDL_EINVAL, "edl_unique: n=%lld out of range", (long long)n);
  EDL_REQUIRE(key_bits >= 1 && key_bits <= 64, EDL_EINVAL, "edl_unique: key_bits=%d not in 1..64", key_bits);
  EDL_REQUIRE(field_bits >= 0 && field_bits <= 64, EDL_EINVAL, "edl_unique: field_bits=%d not in 0..64", field_bits);
  EDL_REQUIRE(seg_offsets && n_unique, EDL_EINVAL, "edl_unique: seg_offsets and n_unique are required");
  EDL_REQUIRE(!field_base || n_fields > 0, EDL_EINVAL, "edl_unique: n_fields must be > 0 with field_base");
  if (n == 0) {
    k_empty_unique<<<1, 1, 0, st>>>(n_unique, seg_offsets);
    EDL_LAUNCH_CHECK("k_empty_unique");
    return EDL_OK;
  }
  EDL_REQUIRE(uniq && order, EDL_EINVAL, "edl_unique: uniq and order are required");
  EDL_REQUIRE(ids, EDL_EINVAL, "edl_unique: ids is null");
  UniqueWs W = carve_unique(workspace, n, key_bits, n_fields, field_bits);
  EDL_REQUIRE(workspace && workspace_bytes >= W.bytes, EDL_EWORKSPACE,
              "edl_unique: workspace %zu < required %zu", workspace_bytes, W.bytes);
  if (sort_keys_32(n, n_dev, field_base, n_fields, key_bits, field_bits))
    return unique_impl<uint32_t>(ids, n, n_dev, field_base, n_fields, key_bits, field_bits, uniq, inverse, counts,
                                 order, seg_offsets, seg_first, n_unique, W, st);
  return unique_impl<uint64_t>(ids, n, n_dev, field_base, n_fields, key_bits, field_bits, uniq, inverse, counts,
                               order, seg_offsets, seg_first, n_unique, W, st);
}

// ------------------------------------------------------- A1 on the owner side of the slot exchange
// The owner's keys arrive as S slots, slot s holding what rank s sent: strictly increasing keys (a stable
// partition of a sorted unique list).  The stable sorted order of their concatenation therefore needs no sort:
// an element's place is its place in its own run plus, for every other run, the number of keys there that go
// before it (smaller key; equal key in a lower slot) -- S - 1 binary searches over L2-resident runs.  One
// kernel replaces edl_slot_compact + the four radix passes of edl_unique + the two edl_remap_index calls: it
// writes the sorted keys and, as `order`, the slot-space rows (bit 30: the gradient row is an fp32 row), so
// the unique kernels behind it emit seg_first in slot space as well.
namespace {
constexpr int kMergeMaxSlots = 64;

constexpr int kMK = 4;   // consecutive keys of a run per thread

// does element v of a run go in front of key k of slot s?  smaller keys do, equal keys when the run's slot is lower
__device__ __forceinline__ bool merge_before(long long v, long long k, bool lower_slot) {
  v &= ~EDL_KEY_FLAG_SINGLE;
  return v < k || (v == k && lower_slot);
}
// number of elements of run[lo0 .. len) in front of k, plus lo0: binary search
__device__ __forceinline__ int merge_count(const long long* __restrict__ run, int lo0, int len, long long k, bool lower_slot) {
  int lo = lo0, hi = len;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (merge_before(run[mid], k, lower_slot)) lo = mid + 1; else hi = mid;
  }
  return lo;
}
// the same when the answer is expected just behind `from` (the count of the thread's previous, smaller key):
// doubling steps, then a binary search inside the last step
__device__ __forceinline__ int merge_advance(const long long* __restrict__ run, int from, int len, long long k, bool lower_slot) {
  if (from >= len || !merge_before(run[from], k, lower_slot)) return from;
  int prev = from, step = 1;           // run[prev] goes in front of k
  while (true) {
    const int nx = prev + step;
    if (nx >= len) return merge_count(run, prev + 1, len, k, lower_slot);
    if (!merge_before(run[nx], k, lower_slot)) return merge_count(run, prev + 1, nx, k, lower_slot);
    prev = nx;
    step <<= 1;
  }
}

// A thread places kMK consecutive keys: for every other run one binary search for its first key, then short walks
// from there for the next ones (keys of a run ascend, so the counts do too) -- a third of the probes of a search
// per key, and the probes, 32 different sectors per warp-wide load, are what this kernel's time is made of.
__global__ void __launch_bounds__(kThreads)
k_slot_merge(const long long* __restrict__ slots, const int32_t* __restrict__ counts, int S, long long stride,
             long long limit, uint64_t* __restrict__ sorted, uint32_t* __restrict__ order, int32_t* __restrict__ n_out) {
  __shared__ int start[kMergeMaxSlots + 1];
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int s = 0; s < S; ++s) {
      start[s] = acc;
      const int c = counts[s];
      acc += (int)(c > limit ? limit : (c < 0 ? 0 : c));
    }
    start[S] = acc;
    if (blockIdx.x == 0) *n_out = acc;
  }
  __syncthreads();
  const int total = start[S];
  for (int i0 = ((int)blockIdx.x * kThreads + threadIdx.x) * kMK; i0 < total; i0 += (int)gridDim.x * kThreads * kMK) {
    long long key[kMK], raw[kMK], src[kMK];
    int slot[kMK], pos[kMK];
    int s = 0;
#pragma unroll
    for (int q = 0; q < kMK; ++q) {
      const int i = i0 + q;
      slot[q] = -1;
      if (i < total) {
        while (start[s + 1] <= i) ++s;    // S <= 64, shared memory
        slot[q] = s;
        pos[q] = i - start[s];
        src[q] = (long long)s * stride + pos[q];
        raw[q] = slots[src[q]];
        key[q] = raw[q] & ~EDL_KEY_FLAG_SINGLE;
      }
    }
    for (int t = 0; t < S; ++t) {
      const long long* run = slots + (long long)t * stride;
      const int len = start[t + 1] - start[t];
      int c = -1;                         // count for the previous key of the same slot, if any
#pragma unroll
      for (int q = 0; q < kMK; ++q) {
        if (slot[q] < 0) continue;
        if (slot[q] == t) { c = -1; continue; }
        const bool lower = t < slot[q];
        if (c < 0 || slot[q] != slot[q - (q > 0)]) c = merge_count(run, 0, len, key[q], lower);
        else c = merge_advance(run, c, len, key[q], lower);
        pos[q] += c;
      }
    }
#pragma unroll
    for (int q = 0; q < kMK; ++q) {
      if (slot[q] < 0) continue;
      sorted[pos[q]] = (uint64_t)key[q];
      order[pos[q]] = (uint32_t)src[q] | (((raw[q] >> 62) & 1) ? (1u << 30) : 0u);
    }
  }
}

struct MergeWs { uint64_t* sorted; uint32_t* tile_heads; size_t bytes; };
MergeWs carve_merge(void* ws, int64_t cap) {
  const int64_t tiles = edl_cdiv(cap > 0 ? cap : 1, kTile);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = edl_align_up(off + bytes, 256); return o; };
  const size_t oS = take((size_t)cap * 8), oT = take((size_t)(tiles + 1) * 4);
  MergeWs r;
  r.sorted = (uint64_t*)((char*)ws + oS); r.tile_heads = (uint32_t*)((char*)ws + oT); r.bytes = off;
  return r;
}
}  // namespace

extern "C" size_t edl_slot_merge_unique_workspace_bytes(int num_slots, int64_t slot_limit) {
  if (num_slots < 1 || slot_limit < 0) return 0;
  return carve_merge(nullptr, (int64_t)num_slots * slot_limit).bytes;
}

extern "C" int edl_slot_merge_unique(const int64_t* keys_slots, const int32_t* counts_dev, int num_slots,
                                     int64_t slot_stride, int64_t slot_limit, int64_t* uniq, int32_t* order,
                                     int32_t* seg_offsets, int32_t* seg_first, int32_t* n_unique, int32_t* n_out_dev,
                                     void* workspace, size_t workspace_bytes, edl_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EDL_REQUIRE(keys_slots && counts_dev && uniq && order && seg_offsets && n_unique && n_out_dev, EDL_EINVAL,
              "edl_slot_merge_unique: null pointer");
  EDL_REQUIRE(num_slots >= 1 && num_slots <= kMergeMaxSlots, EDL_EINVAL, "edl_slot_merge_unique: num_slots=%d out of [1, %d]",
              num_slots, kMergeMaxSlots);
  EDL_REQUIRE(slot_stride > 0 && slot_limit >= 0 && slot_limit <= slot_stride, EDL_EINVAL, "edl_slot_merge_unique: bad slot sizes");
  const int64_t cap = slot_limit * num_slots;
  EDL_REQUIRE(slot_stride * num_slots < (int64_t)(1 << 30) && cap < (int64_t)(1 << 30), EDL_ELIMIT,
              "edl_slot_merge_unique: slots do not fit a 30-bit row index");
  if (cap == 0) {
    k_empty_unique<<<1, 1, 0, st>>>(n_unique, seg_offsets);
    EDL_CUDA(cudaMemsetAsync(n_out_dev, 0, 4, st));
    EDL_LAUNCH_CHECK("k_empty_unique");
    return EDL_OK;
  }
  MergeWs W = carve_merge(workspace, cap);
  EDL_REQUIRE(workspace && workspace_bytes >= W.bytes, EDL_EWORKSPACE, "edl_slot_merge_unique: workspace %zu < required %zu",
              workspace_bytes, W.bytes);
  const int g = (int)std::min<int64_t>(edl_cdiv(cap, kThreads * kMK), EDL_SM_COUNT * 8);
  k_slot_merge<<<g, kThreads, 0, st>>>((const long long*)keys_slots, counts_dev, num_slots, (long long)slot_stride,
                                       (long long)slot_limit, W.sorted, (uint32_t*)order, n_out_dev);
  const int tiles = (int)edl_cdiv(cap, kTile);
  k_head_count<uint64_t><<<tiles, kThreads, 0, st>>>(W.sorted, cap, n_out_dev, W.tile_heads, 0);
  k_scan_tiles<<<1, 1024, 0, st>>>(W.tile_heads, tiles, cap, n_out_dev, n_unique, seg_offsets);
  k_unique_write<uint64_t><<<tiles, kThreads, 0, st>>>(W.sorted, (const uint32_t*)order, cap, n_out_dev, W.tile_heads, uniq,
                                                       seg_offsets, seg_first, (int32_t*)nullptr, 0, (const int64_t*)nullptr);
  EDL_LAUNCH_CHECK("slot merge + unique kernels");
  return EDL_OK;
}

// --------------------------------------------------------------------------------- A2
namespace {
struct PartWs { uint8_t* dig; uint32_t* hist; uint32_t* tot; size_t bytes; };
PartWs carve_part(void* ws, int64_t n) {
  const int64_t tiles = edl_cdiv(n > 0 ? n : 1, kTile);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = edl_align_up(off + bytes, 256); return o; };
  size_t oD = take((size_t)n), oH = take((size_t)256 * tiles * 4), oT = take(256 * 4);
  PartWs r; char* p = (char*)ws;
  r.dig = (uint8_t*)(p + oD); r.hist = (uint32_t*)(p + oH); r.tot = (uint32_t*)(p + oT); r.bytes = off;
  return r;
}
__global__ void k_zero_offsets(int32_t* offsets, int nb) {
  if (threadIdx.x <= nb) offsets[threadIdx.x] = 0;
}
}  // namespace

extern "C" size_t edl_partition_mod_workspace_bytes(int64_t n, int num_shards) {
  (void)num_shards;
  if (n < 0) return 0;
  return carve_part(nullptr, n).bytes;
}

extern "C" int edl_partition_mod(const int64_t* keys, int64_t n, const int32_t* n_dev,
                                 const int64_t* layout, int num_shards, int64_t* keys_out,
                                 int32_t* offsets, int32_t* perm, int32_t* pos, void* workspace,
                                 size_t workspace_bytes, edl_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EDL_REQUIRE(n >= 0 && n < (int64_t)0x7fff0000, EDL_EINVAL, "edl_partition_mod: n=%lld out of range", (long long)n);
  EDL_REQUIRE(num_shards >= 1 && num_shards <= 255, EDL_EINVAL,
              "edl_partition_mod: num_shards=%d not in 1..255", num_shards);
  EDL_REQUIRE(offsets, EDL_EINVAL, "edl_partition_mod: offsets is null");
  if (n == 0) {
    k_zero_offsets<<<1, 256, 0, st>>>(offsets, num_shards);
    EDL_LAUNCH_CHECK("k_zero_offsets");
    return EDL_OK;
  }
  EDL_REQUIRE(keys && keys_out && perm, EDL_EINVAL, "edl_partition_mod: null keys / keys_out / perm");
  PartWs W = carve_part(workspace, n);
  EDL_REQUIRE(workspace && workspace_bytes >= W.bytes, EDL_EWORKSPACE,
              "edl_partition_mod: workspace %zu < required %zu", workspace_bytes, W.bytes);
  const int tiles = (int)edl_cdiv(n, kTile);
  int g = (int)std::min<int64_t>(edl_cdiv(n, kThreads), EDL_SM_COUNT * 8);
  k_shard_digits<<<g, kThreads, 0, st>>>(keys, n, n_dev, layout, (unsigned)num_shards, W.dig);
  KeyIndexLoader ld{keys};
  ArrayDigit dg{W.dig};
  const Seg flat{0, 0, 0, 0};
  k_hist<int64_t><<<tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, tiles, W.hist, flat);
  k_rowscan<<<256, kThreads, 0, st>>>(W.hist, tiles, W.tot, 0);
  k_scatter<int64_t><<<tiles, kThreads, 0, st>>>(ld, dg, n, n_dev, tiles, W.hist, W.tot, keys_out,
                                                 (uint32_t*)perm, pos, offsets, num_shards, flat);
  EDL_LAUNCH_CHECK("partition kernels");
  return EDL_OK;
}

// --------------------------------------------------------------- A1 without any allocation
// edl_unique_static: the same dedup as edl_unique, but the scratch it needs lives in a fixed
// array that is part of the library image instead of a caller-provided workspace.
//
// The array below is DEVICE memory.  Inside host functions its name is only a handle the
// CUDA runtime understands -- it is not a device address and must never be handed to a kernel
// as a pointer.  The real address is asked for once per device with cudaGetSymbolAddress
// (with that device current) and cached in g_static_addr[].

static __device__ __align__(256) unsigned char g_unique_scratch[EDL_UNIQUE_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_addr[kMaxDevices] = {nullptr};

// Resolves the device address of g_unique_scratch on the current device (first call only).
int static_scratch(void** out) {
  int dev = -1;
  EDL_CUDA(cudaGetDevice(&dev));
  EDL_REQUIRE(dev >= 0 && dev < kMaxDevices, EDL_ELIMIT, "edl_unique_static: device ordinal %d beyond %d",
              dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[dev]) {
    void* p = nullptr;
    EDL_CUDA(cudaGetSymbolAddress(&p, g_unique_scratch));
    EDL_REQUIRE(p != nullptr, EDL_ECUDA, "edl_unique_static: cudaGetSymbolAddress returned a null address");
    g_static_addr[dev] = p;
  }
  *out = g_static_addr[dev];
  return EDL_OK;
}
}  // namespace

extern "C" int edl_unique_static(const int64_t* ids, int64_t n, const int32_t* n_dev,
                                 const int64_t* field_base, int n_fields, int key_bits,
                                 int field_bits, int64_t* uniq, int32_t* inverse, int32_t* counts, int32_t* order,
                                 int32_t* seg_offsets, int32_t* seg_first, int32_t* n_unique,
                                 edl_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EDL_REQUIRE(n >= 0, EDL_EINVAL, "edl_unique_static: n=%lld negative", (long long)n);
  EDL_REQUIRE(n <= (int64_t)EDL_UNIQUE_STATIC_MAX_IDS, EDL_ELIMIT,
              "edl_unique_static: n=%lld exceeds the built-in scratch limit of %d ids; use edl_unique "
              "with a caller-provided workspace for larger batches",
              (long long)n, EDL_UNIQUE_STATIC_MAX_IDS);
  EDL_REQUIRE(key_bits >= 1 && key_bits <= 64, EDL_EINVAL, "edl_unique_static: key_bits=%d not in 1..64", key_bits);
  EDL_REQUIRE(seg_offsets && n_unique, EDL_EINVAL, "edl_unique_static: seg_offsets and n_unique are required");
  EDL_REQUIRE(!field_base || n_fields > 0, EDL_EINVAL, "edl_unique_static: n_fields must be > 0 with field_base");
  if (n == 0) {
    k_empty_unique<<<1, 1, 0, st>>>(n_unique, seg_offsets);
    EDL_LAUNCH_CHECK("k_empty_unique");
    return EDL_OK;
  }
  EDL_REQUIRE(ids && uniq && order, EDL_EINVAL, "edl_unique_static: ids, uniq and order are required");
  void* scratch = nullptr;
  int rc = static_scratch(&scratch);
  if (rc) return rc;
  EDL_REQUIRE(field_bits >= 0 && field_bits <= 64, EDL_EINVAL, "edl_unique_static: field_bits=%d not in 0..64", field_bits);
  UniqueWs W = carve_unique(scratch, n, key_bits, n_fields, field_bits);
  EDL_REQUIRE(W.bytes <= (size_t)EDL_UNIQUE_STATIC_SCRATCH_BYTES, EDL_ELIMIT,
              "edl_unique_static: n=%lld at key_bits=%d needs %zu scratch bytes, the built-in array has %d",
              (long long)n, key_bits, W.bytes, EDL_UNIQUE_STATIC_SCRATCH_BYTES);
  // Nothing in the scratch has to be zero when this call starts: every histogram cell, digit total
  // and tile count is written by this call before this call reads it, and the few words the fused
  // sort kernel needs at zero (its ticket and per-field totals) are cleared by this call on `stream`.
  if (sort_keys_32(n, n_dev, field_base, n_fields, key_bits, field_bits))
    return unique_impl<uint32_t>(ids, n, n_dev, field_base, n_fields, key_bits, field_bits, uniq, inverse, counts,
                                 order, seg_offsets, seg_first, n_unique, W, st);
  return unique_impl<uint64_t>(ids, n, n_dev, field_base, n_fields, key_bits, field_bits, uniq, inverse, counts,
                               order, seg_offsets, seg_first, n_unique, W, st);
}
