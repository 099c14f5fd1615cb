// A7 without a host round trip: exchange rounds over fixed per-pair slots.
//
// The packed all-to-all-v of peer.cu needs the count matrix on the host (receive offsets depend on
// what every other rank sends), which costs the step its one host synchronisation.  Here every
// receive buffer is cut into `world` slots of `slot_stride` rows, slot s written only by rank s,
// so a sender needs nothing but its own partition offsets -- which stay on the device.  The row
// counts travel with the data (one int32 per pair).  With no size known to the host, the whole
// sharded step is a fixed launch sequence and can be replayed from a CUDA graph.
#include "edl_common.cuh"

namespace {

constexpr int kT = 256;
constexpr int kMaxSlots = 64;

__device__ __forceinline__ int slot_of(const int* __restrict__ offs, int S, int p) {
  int d = 0;
#pragma unroll 1
  for (int s = 1; s < S; ++s) d += (p >= offs[s]) ? 1 : 0;   // offsets ascending: count the starts passed
  return d;
}

// rows[p] (partition order, group d = offsets[d]..offsets[d+1]) -> dst_bases[d] + (p - offsets[d]) rows.
// One thread per 8-byte word of a row.  Thread d of block 0 also publishes group d's row count.
__global__ void __launch_bounds__(kT)
k_slot_send(const unsigned long long* __restrict__ rows, int words, const int32_t* __restrict__ offsets, int S,
            const long long* __restrict__ dst_bases, long long slot_limit,
            const long long* __restrict__ count_ptrs, int32_t* __restrict__ err,
            const int32_t* __restrict__ flag_seg, const int32_t* __restrict__ flag_perm) {
  __shared__ int offs[kMaxSlots + 1];
  for (int i = threadIdx.x; i <= S; i += kT) offs[i] = offsets[i];
  __syncthreads();
  if (blockIdx.x == 0 && (int)threadIdx.x < S) {
    const int d = threadIdx.x;
    const int c = offs[d + 1] - offs[d];
    if (c > slot_limit) atomicOr(err, 1);
    if (count_ptrs) *(int32_t*)count_ptrs[d] = (int32_t)(c > slot_limit ? slot_limit : c);
  }
  const long long total = (long long)offs[S] * words;
  for (long long i = (long long)blockIdx.x * kT + threadIdx.x; i < total; i += (long long)gridDim.x * kT) {
    const int p = (int)(i / words), w = (int)(i - (long long)p * words);
    const int d = slot_of(offs, S, p);
    const int k = p - offs[d];
    if (k >= slot_limit) continue;
    unsigned long long v = rows[i];
    if (flag_seg) {   // keys (one word per row): bit 62 marks an id with exactly one local gradient row
      const int u = flag_perm[p];
      if (flag_seg[u + 1] - flag_seg[u] == 1) v |= (unsigned long long)EDL_KEY_FLAG_SINGLE;
    }
    ((unsigned long long*)dst_bases[d])[(long long)k * words + w] = v;
  }
}

// out[u] = slot-space row of partition position pos[u]: d * slot_stride + (pos[u] - offsets[d])
__global__ void __launch_bounds__(kT)
k_slot_positions(const int32_t* __restrict__ pos, long long n_max, const int32_t* __restrict__ n_dev,
                 const int32_t* __restrict__ offsets, int S, long long slot_stride, int32_t* __restrict__ out) {
  __shared__ int offs[kMaxSlots + 1];
  for (int i = threadIdx.x; i <= S; i += kT) offs[i] = offsets[i];
  __syncthreads();
  const long long n = n_dev ? min((long long)*n_dev, n_max) : n_max;
  for (long long u = (long long)blockIdx.x * kT + threadIdx.x; u < n; u += (long long)gridDim.x * kT) {
    const int p = pos[u];
    const int d = slot_of(offs, S, p);
    long long k = p - offs[d];
    if (k >= slot_stride) k = slot_stride - 1;   // a cut group (flagged by the sender): stay inside the slot
    out[u] = (int32_t)((long long)d * slot_stride + k);
  }
}

// The owner's view: S slots of keys with counts[s] valid rows each -> one contiguous key array in
// slot (= peer rank) order, the slot-space row of each, and the total.
__global__ void __launch_bounds__(kT)
k_slot_compact(const long long* __restrict__ keys_slots, const int32_t* __restrict__ counts, int S,
               long long slot_stride, long long slot_limit, long long* __restrict__ keys_out,
               int32_t* __restrict__ map_out, int32_t* __restrict__ n_out) {
  __shared__ int start[kMaxSlots + 1];
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int s = 0; s < S; ++s) {
      start[s] = acc;
      const int c = counts[s];
      acc += (int)(c > slot_limit ? slot_limit : (c < 0 ? 0 : c));
    }
    start[S] = acc;
    if (blockIdx.x == 0) *n_out = acc;
  }
  __syncthreads();
  const long long total = start[S];
  for (long long i = (long long)blockIdx.x * kT + threadIdx.x; i < total; i += (long long)gridDim.x * kT) {
    const int s = slot_of(start, S, (int)i);
    const long long src = (long long)s * slot_stride + (i - start[s]);
    const long long k = keys_slots[src];
    keys_out[i] = k & ~EDL_KEY_FLAG_SINGLE;
    map_out[i] = (int32_t)src | (((k >> 62) & 1) ? (1 << 30) : 0);   // bit 30: the gradient row is an fp32 row
  }
}

__global__ void __launch_bounds__(kT)
k_remap(const int32_t* __restrict__ index, const int32_t* __restrict__ map, long long n_max,
        const int32_t* __restrict__ n_dev, int32_t* __restrict__ out) {
  const long long n = n_dev ? min((long long)*n_dev, n_max) : n_max;
  for (long long i = (long long)blockIdx.x * kT + threadIdx.x; i < n; i += (long long)gridDim.x * kT)
    out[i] = map[index[i]];
}

// k_signal_wait of peer.cu with the round number kept on the device (incremented here), so that a
// captured graph publishes a fresh value on every replay.
__global__ void k_signal_wait_counter(long long* const* peer_flags, volatile long long* own_flags, int me,
                                      int world, long long* counter) {
  __shared__ long long value_s;
  if (threadIdx.x == 0) { value_s = *counter + 1; *counter = value_s; }
  __syncthreads();
  const long long value = value_s;
  const int r = threadIdx.x;
  if (r >= world) return;
  __threadfence_system();
  long long* dst = peer_flags[r] + me;
  asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(dst), "l"(value) : "memory");
  long long seen = 0;
  for (long long spins = 0; spins < (1ll << 27); ++spins) {
    asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(seen) : "l"(own_flags + r) : "memory");
    if (seen >= value) break;
  }
  if (seen < value) own_flags[world] = -1;
  __threadfence_system();
}

inline unsigned grid_for(long long items) {
  const long long want = (items + kT - 1) / kT;
  const long long cap = (long long)EDL_SM_COUNT * 8;
  return (unsigned)(want < 1 ? 1 : (want > cap ? cap : want));
}

}  // namespace

extern "C" int edl_slot_send(const void* rows, int row_bytes, const int32_t* offsets_dev, int num_slots,
                             const int64_t* dst_bases_dev, int64_t slot_limit, const int64_t* count_ptrs_dev,
                             int32_t* err_dev, int64_t n_max, const int32_t* flag_seg_offsets, const int32_t* flag_perm,
                             edl_stream_t stream) {
  EDL_REQUIRE(num_slots >= 1 && num_slots <= kMaxSlots, EDL_EINVAL, "edl_slot_send: num_slots=%d out of [1, %d]",
              num_slots, kMaxSlots);
  EDL_REQUIRE(row_bytes > 0 && row_bytes % 8 == 0, EDL_EINVAL, "edl_slot_send: row_bytes=%d must be a positive multiple of 8", row_bytes);
  EDL_REQUIRE(offsets_dev && dst_bases_dev && err_dev, EDL_EINVAL, "edl_slot_send: null pointer");
  EDL_REQUIRE(n_max >= 0 && slot_limit >= 0, EDL_EINVAL, "edl_slot_send: negative size");
  EDL_REQUIRE(rows || n_max == 0, EDL_EINVAL, "edl_slot_send: rows is null");
  EDL_REQUIRE((((uintptr_t)rows) & 7) == 0, EDL_EINVAL, "edl_slot_send: rows must be 8-byte aligned");
  EDL_REQUIRE(!flag_seg_offsets || (flag_perm && row_bytes == 8), EDL_EINVAL,
              "edl_slot_send: the single-row flag applies to 8-byte keys and needs flag_perm");
  k_slot_send<<<grid_for(n_max * (row_bytes / 8)), kT, 0, (cudaStream_t)stream>>>(
      (const unsigned long long*)rows, row_bytes / 8, offsets_dev, num_slots, (const long long*)dst_bases_dev,
      (long long)slot_limit, (const long long*)count_ptrs_dev, err_dev, flag_seg_offsets, flag_perm);
  EDL_LAUNCH_CHECK("k_slot_send");
  return EDL_OK;
}

extern "C" int edl_slot_positions(const int32_t* pos, int64_t n_max, const int32_t* n_dev, const int32_t* offsets_dev,
                                  int num_slots, int64_t slot_stride, int32_t* out, edl_stream_t stream) {
  if (n_max == 0) return EDL_OK;
  EDL_REQUIRE(num_slots >= 1 && num_slots <= kMaxSlots, EDL_EINVAL, "edl_slot_positions: num_slots=%d out of range", num_slots);
  EDL_REQUIRE(pos && offsets_dev && out && n_max > 0, EDL_EINVAL, "edl_slot_positions: bad argument");
  EDL_REQUIRE(slot_stride > 0 && slot_stride * num_slots < (int64_t)0x7fffffff, EDL_ELIMIT,
              "edl_slot_positions: %d slots x %lld rows do not fit an int32 row index", num_slots, (long long)slot_stride);
  k_slot_positions<<<grid_for(n_max), kT, 0, (cudaStream_t)stream>>>(pos, (long long)n_max, n_dev, offsets_dev, num_slots,
                                                                    (long long)slot_stride, out);
  EDL_LAUNCH_CHECK("k_slot_positions");
  return EDL_OK;
}

extern "C" int edl_slot_compact(const int64_t* keys_slots, const int32_t* counts_dev, int num_slots, int64_t slot_stride,
                                int64_t slot_limit, int64_t* keys_out, int32_t* map_out, int32_t* n_out_dev,
                                edl_stream_t stream) {
  EDL_REQUIRE(num_slots >= 1 && num_slots <= kMaxSlots, EDL_EINVAL, "edl_slot_compact: num_slots=%d out of range", num_slots);
  EDL_REQUIRE(keys_slots && counts_dev && keys_out && map_out && n_out_dev, EDL_EINVAL, "edl_slot_compact: null pointer");
  EDL_REQUIRE(slot_stride > 0 && slot_limit >= 0 && slot_limit <= slot_stride, EDL_EINVAL, "edl_slot_compact: bad slot sizes");
  EDL_REQUIRE(slot_stride * num_slots < (int64_t)(1 << 30) && slot_limit * num_slots < (int64_t)(1 << 30), EDL_ELIMIT,
              "edl_slot_compact: slots do not fit a 30-bit row index");
  k_slot_compact<<<grid_for(slot_limit * num_slots), kT, 0, (cudaStream_t)stream>>>(
      (const long long*)keys_slots, counts_dev, num_slots, (long long)slot_stride, (long long)slot_limit,
      (long long*)keys_out, map_out, n_out_dev);
  EDL_LAUNCH_CHECK("k_slot_compact");
  return EDL_OK;
}

extern "C" int edl_remap_index(const int32_t* index, const int32_t* map, int64_t n_max, const int32_t* n_dev,
                               int32_t* out, edl_stream_t stream) {
  if (n_max == 0) return EDL_OK;
  EDL_REQUIRE(index && map && out && n_max > 0, EDL_EINVAL, "edl_remap_index: bad argument");
  k_remap<<<grid_for(n_max), kT, 0, (cudaStream_t)stream>>>(index, map, (long long)n_max, n_dev, out);
  EDL_LAUNCH_CHECK("k_remap");
  return EDL_OK;
}

extern "C" int edl_peer_signal_wait_counter(const int64_t* peer_flag_ptrs_dev, int64_t* own_flags, int me, int world,
                                            int64_t* counter_dev, edl_stream_t stream) {
  EDL_REQUIRE(peer_flag_ptrs_dev && own_flags && counter_dev, EDL_EINVAL, "edl_peer_signal_wait_counter: null pointer");
  EDL_REQUIRE(world >= 1 && world <= 64 && me >= 0 && me < world, EDL_EINVAL,
              "edl_peer_signal_wait_counter: bad rank / world");
  k_signal_wait_counter<<<1, 64, 0, (cudaStream_t)stream>>>((long long* const*)peer_flag_ptrs_dev,
                                                             (volatile long long*)own_flags, me, world,
                                                             (long long*)counter_dev);
  EDL_LAUNCH_CHECK("k_signal_wait_counter");
  return EDL_OK;
}
