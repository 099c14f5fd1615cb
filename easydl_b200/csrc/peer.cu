// A7 over NVLink peer memory: set-up helpers for exchange buffers that other ranks of the box
// write into directly (CUDA IPC), and the peer copy itself.  These are the only entry points of
// the library that allocate: a CUDA IPC handle can only be exported for a whole cudaMalloc
// allocation, so the exchange buffers cannot come from a framework's sub-allocating pool.
// Set-up time only; nothing here runs inside a training step except edl_peer_copy.
#include <string.h>

#include "edl_common.cuh"

extern "C" int edl_peer_alloc(void** ptr, size_t bytes) {
  EDL_REQUIRE(ptr && bytes > 0, EDL_EINVAL, "edl_peer_alloc: bad argument");
  EDL_CUDA(cudaMalloc(ptr, bytes));
  EDL_CUDA(cudaMemset(*ptr, 0, bytes));
  return EDL_OK;
}

extern "C" int edl_peer_free(void* ptr) {
  if (ptr) EDL_CUDA(cudaFree(ptr));
  return EDL_OK;
}

extern "C" int edl_ipc_export(void* ptr, unsigned char* handle64) {
  EDL_REQUIRE(ptr && handle64, EDL_EINVAL, "edl_ipc_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle is 64 bytes");
  cudaIpcMemHandle_t h;
  EDL_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, 64);
  return EDL_OK;
}

extern "C" int edl_ipc_import(const unsigned char* handle64, void** ptr) {
  EDL_REQUIRE(ptr && handle64, EDL_EINVAL, "edl_ipc_import: null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  EDL_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return EDL_OK;
}

extern "C" int edl_ipc_close(void* ptr) {
  if (ptr) EDL_CUDA(cudaIpcCloseMemHandle(ptr));
  return EDL_OK;
}

// dst may live on a peer GPU (an edl_ipc_import'ed pointer): the copy runs over NVLink
extern "C" int edl_peer_copy(void* dst, const void* src, size_t bytes, edl_stream_t stream) {
  if (bytes == 0) return EDL_OK;
  EDL_REQUIRE(dst && src, EDL_EINVAL, "edl_peer_copy: null pointer");
  EDL_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return EDL_OK;
}

// Cross-GPU fence for one exchange round, as ONE kernel per rank: lane r publishes `value` into
// peer r's flag slot for this rank (after everything this stream wrote before), then waits until
// peer r's `value` has arrived in this rank's own slots.  Every rank signals before it waits, so
// ranks never wait on each other's progress through this kernel -- one rank per GPU only (see
// B200_PROFILING.md on waiting kernels sharing a GPU).  A bounded spin: on timeout the kernel
// stores -1 in own_flags[world] and returns instead of hanging the box.
namespace {
__global__ void k_signal_wait(long long* const* peer_flags, volatile long long* own_flags, int me, int world,
                              long long value) {
  const int r = threadIdx.x;
  if (r >= world) return;
  __threadfence_system();
  long long* dst = peer_flags[r] + me;
  asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(dst), "l"(value) : "memory");
  long long seen = 0;
  for (long long spins = 0; spins < (1ll << 27); ++spins) {   // ~2 minutes of polling
    asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(seen) : "l"(own_flags + r) : "memory");
    if (seen >= value) break;
  }
  if (seen < value) own_flags[world] = -1;
  __threadfence_system();
}
}  // namespace

extern "C" int edl_peer_signal_wait(const int64_t* peer_flag_ptrs_dev, int64_t* own_flags, int me, int world,
                                    int64_t value, edl_stream_t stream) {
  EDL_REQUIRE(peer_flag_ptrs_dev && own_flags, EDL_EINVAL, "edl_peer_signal_wait: null pointer");
  EDL_REQUIRE(world >= 1 && world <= 64 && me >= 0 && me < world, EDL_EINVAL, "edl_peer_signal_wait: bad rank / world");
  k_signal_wait<<<1, 64, 0, (cudaStream_t)stream>>>((long long* const*)peer_flag_ptrs_dev,
                                                     (volatile long long*)own_flags, me, world, (long long)value);
  EDL_LAUNCH_CHECK("k_signal_wait");
  return EDL_OK;
}
