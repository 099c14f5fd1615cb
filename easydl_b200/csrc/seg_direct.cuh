// Register-path segmented reduce + optimizer: one lane group (dim/4 lanes) per unique id, rows
// read straight from global memory with kUnrollD loads in flight, thousands of groups resident
// per SM.  Hot ids (more than kLongThr gradient rows) are cut into kChunk-row chunks that are
// summed by separate lane groups (k_seg_chunk) -- balanced under any Zipf skew and deterministic.  Included by update.cu.
#pragma once
#include <mutex>
#include "seg_common.cuh"
#include "seg_stream.cuh"

namespace {

constexpr int kLongThrMin = 2;  // smallest allowed hot-id threshold (sizes the hot-id lists; knob long_thr, default 16)
constexpr int kChunk = 64;     // gradient rows per chunk
constexpr int kUnrollD = 4;    // independent row loads in flight per lane group (short path)
constexpr int kUnrollC = 2;    // same, chunk path


// kMode: 0 every segment; 1 only ids that occur once; 2 only ids that occur more than once
template <int kG, int kC, typename GT, typename Epi, int kMinBlocks = 4, int kUnroll = kUnrollD, int kMode = 0, int kBlock = kThreads,
          int kBatch = 1>
__global__ void __launch_bounds__(kBlock, kMinBlocks)
k_seg_direct(const unsigned char* __restrict__ grad, const unsigned char* __restrict__ grad32, int D4,
             const int32_t* __restrict__ order,
             const int32_t* __restrict__ seg, const int32_t* __restrict__ seg_first,
             const int32_t* __restrict__ n_unique,
             const int64_t* __restrict__ uniq, const int64_t* __restrict__ layout, Epi epi,
             int32_t* __restrict__ counters, int32_t* __restrict__ long_list,
             int32_t* __restrict__ chunk_list, int kLongThr) {
  __shared__ EdlLayoutSmem L;
  typedef Row<GT> R;
  edl_layout_load(&L, layout);
  __syncthreads();
  epi.prepare();
  const int U = *n_unique;
  const int gl = threadIdx.x % kG;
  // Each lane group takes kBatch ids, one CTA-width of groups apart.  The index loads (segment
  // bounds, first row, key) of ALL of them are issued before the first id's rows are touched, so the
  // first memory hop of ids 2..kBatch is hidden behind the work on the id before.
  constexpr int kGroups = kBlock / kG;
  const int ub = (int)blockIdx.x * (kGroups * kBatch) + (int)threadIdx.x / kG;
  int s0a[kBatch], s1a[kBatch], j0a[kBatch];
  long long toka[kBatch];
#pragma unroll
  for (int b = 0; b < kBatch; ++b) {
    const int ux = ub + b * kGroups;
    if (ux < U) {
      s0a[b] = seg[ux]; s1a[b] = seg[ux + 1];
      j0a[b] = seg_first ? seg_first[ux] : -1;
      toka[b] = epi.token(uniq, ux);
    }
  }
#pragma unroll
  for (int b = 0; b < kBatch; ++b) {
    const int u = ub + b * kGroups;
    if (u >= U) break;
    const int s0 = s0a[b], s1 = s1a[b], j0 = j0a[b];
    const long long tok = toka[b];
    [&]() {
  const int cnt = s1 - s0;
  if (kMode == 1 && cnt != 1) return;
  if (kMode == 2 && cnt == 1) return;
  if (kMode != 1 && cnt > kLongThr) {
    if (!counters) return;   // k_seg_classify listed the hot ids already (forked launch)
    // hot id: register its chunks (one lane claims the slots, the group fills them)
    const int nch = (cnt + kChunk - 1) / kChunk;
    int base = 0, li = 0;
    if (gl == 0) {
      base = atomicAdd(&counters[1], nch);
      li = atomicAdd(&counters[0], 1);
      long_list[li * 4 + 0] = u; long_list[li * 4 + 1] = base; long_list[li * 4 + 2] = nch;
      long_list[li * 4 + 3] = nch > 1 ? atomicAdd(&counters[3], nch) : 0;   // first partial row
    }
    // broadcast from the group's first lane; only this group's lanes take part
    const unsigned lane = threadIdx.x & 31u;
    const unsigned gmask = kG == 32 ? 0xffffffffu : (((1u << kG) - 1u) << (lane / kG * kG));
    base = __shfl_sync(gmask, base, lane / kG * kG);
    li = __shfl_sync(gmask, li, lane / kG * kG);
    for (int c = gl; c < nch; c += kG) {   // begin / end stored here: the chunk kernel needs one hop less
      const int b = s0 + c * kChunk;
      int4 ent = make_int4(b, min(s1, b + kChunk), li, c);
      ((int4*)chunk_list)[base + c] = ent;
    }
    return;
  }
  const long long row = epi.row_from(&L, tok);
  if (row < 0) return;   // a write epilogue dropped the row (beyond its slot)
  float4 p[kC], a0[kC], a1[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) {
    const int col = gl + c * kG;
    if (col < D4) epi.load(row, D4, col, &p[c], &a0[c], &a1[c]);   // in flight during the reduce
  }
  if (kMode != 2 && cnt == 1) {
    const int j = j0 >= 0 ? j0 : order[s0];
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int col = gl + c * kG;
      if (col < D4) {
        typename R::V v = R::row(grad, grad32, j, D4, col);
        if constexpr (Epi::kMixed) {
          epi.apply_single(row, D4, col, R::f32(v));   // one gradient row: it travels as it is
        } else {
          double s[4] = {0.0, 0.0, 0.0, 0.0};
          if (Epi::kOutF64) R::add(s, v);
          epi.apply(row, D4, col, R::f32(v), s, p[c], a0[c], a1[c]);
        }
      }
    }
  } else if (kMode != 1) {
    double acc[kC][4];
#pragma unroll
    for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
    for (int i = s0; i < s1; i += kUnroll) {
      typename R::V v[kUnroll][kC];
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        if (i + k < s1) {
          const int j = order[i + k];
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) v[k][c] = R::row(grad, grad32, j, D4, col);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kUnroll; ++k) {
        if (i + k < s1) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) R::add(acc[c], v[k][c]);
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int col = gl + c * kG;
      if (col < D4) {
        float4 g = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                               __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
        epi.apply(row, D4, col, g, acc[c], p[c], a0[c], a1[c]);
      }
    }
  }
  if (gl == 0) epi.touch(row);
    }();
  }
}

// Forked launch only: the hot-id lists made ahead of k_seg_direct (one thread per id), so that the
// chunk kernels do not have to wait for it.  Same list layout as the in-kernel registration.
__global__ void __launch_bounds__(kThreads)
k_seg_classify(const int32_t* __restrict__ seg, const int32_t* __restrict__ n_unique,
               int32_t* __restrict__ counters, int32_t* __restrict__ long_list,
               int32_t* __restrict__ chunk_list, int kLongThr) {
  const int U = *n_unique;
  const int u = (int)blockIdx.x * kThreads + (int)threadIdx.x;
  if (u >= U) return;
  const int s0 = seg[u], s1 = seg[u + 1];
  if (s1 - s0 <= kLongThr) return;
  const int nch = (s1 - s0 + kChunk - 1) / kChunk;
  const int base = atomicAdd(&counters[1], nch);
  const int li = atomicAdd(&counters[0], 1);
  const int pbase = nch > 1 ? atomicAdd(&counters[3], nch) : 0;   // partial rows only where they are combined
  ((int4*)long_list)[li] = make_int4(u, base, nch, pbase);
  for (int c = 0; c < nch; ++c) {
    const int b = s0 + c * kChunk;
    ((int4*)chunk_list)[base + c] = make_int4(b, min(s1, b + kChunk), li, c);
  }
}

// Helper stream of the forked launch, one per device, made on the first call on that device (which
// therefore must not be inside a stream capture).  The mutex keeps one call's record/wait pairs
// on the shared events together when several host threads drive one device.
struct ForkRes { cudaStream_t side; cudaEvent_t fork, join; bool ready; };
inline std::mutex& fork_mutex() {   // one for every instantiation of launch_direct
  static std::mutex mu;
  return mu;
}
inline int fork_res(ForkRes** out) {
  static ForkRes res[64];
  int dev = 0;
  EDL_CUDA(cudaGetDevice(&dev));
  EDL_REQUIRE(dev >= 0 && dev < 64, EDL_ELIMIT, "segreduce: device index %d out of range", dev);
  ForkRes& r = res[dev];
  if (!r.ready) {
    int lo = 0, hi = 0;
    EDL_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    EDL_CUDA(cudaStreamCreateWithPriority(&r.side, cudaStreamNonBlocking, hi));
    EDL_CUDA(cudaEventCreateWithFlags(&r.fork, cudaEventDisableTiming));
    EDL_CUDA(cudaEventCreateWithFlags(&r.join, cudaEventDisableTiming));
    r.ready = true;
  }
  *out = &r;
  return EDL_OK;
}

// One lane group per chunk (<= kChunk rows) of a hot id.  An id that fits one chunk is finished
// here (sum + epilogue).  Longer ids leave a float64 partial row per chunk in scratch for
// k_seg_long_final.
// occupancy target by row width: 32 registers only pay off while one float4 per lane is live
template <int kG, int kC, typename GT, typename Epi>
__global__ void __launch_bounds__(kThreads, (kC == 1 ? 8 : (kC == 2 ? 4 : 2)))
k_seg_chunk(const unsigned char* __restrict__ grad, const unsigned char* __restrict__ grad32, int D4,
            const int32_t* __restrict__ order,
            const int32_t* __restrict__ seg, const int32_t* __restrict__ counters,
            const int32_t* __restrict__ chunk_list, const int32_t* __restrict__ long_list,
            double* __restrict__ partial, const int64_t* __restrict__ uniq,
            const int64_t* __restrict__ layout, Epi epi) {
  __shared__ EdlLayoutSmem L;
  typedef Row<GT> R;
  edl_layout_load(&L, layout);
  __syncthreads();
  epi.prepare();
  const int n_chunks = counters[1];
  const int gl = threadIdx.x % kG;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned gmask = kG == 32 ? 0xffffffffu : (((1u << kG) - 1u) << (lane / kG * kG));
  const int groups = (int)((int64_t)gridDim.x * kThreads / kG);
  const int D = D4 * 4;
  for (int ci = (int)(((int64_t)blockIdx.x * kThreads + threadIdx.x) / kG); ci < n_chunks; ci += groups) {
    const int4 ent = ((const int4*)chunk_list)[ci];
    const int b = ent.x, e = ent.y, li = ent.z;
    const int4 le = ((const int4*)long_list)[li];   // needed after the sum only
    const int u = le.x, nch = le.z;
    double acc[kC][4];
#pragma unroll
    for (int c = 0; c < kC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
    for (int i = b; i < e; i += kUnrollC) {
      typename R::V v[kUnrollC][kC];
#pragma unroll
      for (int k = 0; k < kUnrollC; ++k) {
        if (i + k < e) {
          const int j = order[i + k];
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) v[k][c] = R::row(grad, grad32, j, D4, col);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kUnrollC; ++k) {
        if (i + k < e) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) R::add(acc[c], v[k][c]);
          }
        }
      }
    }
    if (nch > 1) {
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        const int col = gl + c * kG;
        if (col < D4) {
          double2* dst = (double2*)(partial + (size_t)(le.w + ent.w) * D + col * 4);
          dst[0] = make_double2(acc[c][0], acc[c][1]);
          dst[1] = make_double2(acc[c][2], acc[c][3]);
        }
      }
      continue;   // k_seg_long_final adds the partial rows of this id in chunk order
    }
    const long long row = epi.row_of(&L, uniq, u);
    if (row < 0) continue;
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int col = gl + c * kG;
      if (col < D4) {
        float4 p = make_float4(0, 0, 0, 0), a0 = p, a1 = p;
        epi.load(row, D4, col, &p, &a0, &a1);
        float4 g = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                               __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
        epi.apply(row, D4, col, g, acc[c], p, a0, a1);
      }
    }
    if (gl == 0) epi.touch(row);
  }
}

// one lane group per hot id with more than one chunk: partial rows added in chunk order, epilogue
template <int kG, int kC, typename Epi>
__global__ void __launch_bounds__(kThreads, 4)
k_seg_long_final(int D4, const int32_t* __restrict__ counters, const int32_t* __restrict__ long_list,
                 const double* __restrict__ partial, const int64_t* __restrict__ uniq,
                 const int64_t* __restrict__ layout, Epi epi) {
  __shared__ EdlLayoutSmem L;
  edl_layout_load(&L, layout);
  __syncthreads();
  epi.prepare();
  const int n_long = counters[0];
  const int gl = threadIdx.x % kG;
  const int groups = (int)((int64_t)gridDim.x * kThreads / kG);
  const int D = D4 * 4;
  for (int li = (int)(((int64_t)blockIdx.x * kThreads + threadIdx.x) / kG); li < n_long; li += groups) {
    const int4 le = ((const int4*)long_list)[li];
    const int u = le.x, base = le.w, nch = le.z;   // base: first partial row
    if (nch <= 1) continue;   // finished inside k_seg_chunk
    const long long row = epi.row_of(&L, uniq, u);
    if (row < 0) continue;
    float4 p[kC], a0[kC], a1[kC];
    double acc[kC][4];
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
      const int col = gl + c * kG;
      if (col < D4) epi.load(row, D4, col, &p[c], &a0[c], &a1[c]);
    }
    for (int k0 = 0; k0 < nch; k0 += 4) {   // 4 partial rows in flight, added in chunk order
      double2 x[4][kC], y[4][kC];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (k0 + k < nch) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) {
              const double2* src = (const double2*)(partial + (size_t)(base + k0 + k) * D + col * 4);
              x[k][c] = __ldcg(src); y[k][c] = __ldcg(src + 1);
            }
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (k0 + k < nch) {
#pragma unroll
          for (int c = 0; c < kC; ++c) {
            const int col = gl + c * kG;
            if (col < D4) {
              acc[c][0] += x[k][c].x; acc[c][1] += x[k][c].y; acc[c][2] += y[k][c].x; acc[c][3] += y[k][c].y;
            }
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int col = gl + c * kG;
      if (col < D4) {
        float4 g = make_float4(__double2float_rn(acc[c][0]), __double2float_rn(acc[c][1]),
                               __double2float_rn(acc[c][2]), __double2float_rn(acc[c][3]));
        epi.apply(row, D4, col, g, acc[c], p[c], a0[c], a1[c]);
      }
    }
    if (gl == 0) epi.touch(row);
  }
}

// One call = memset of the counters (update.cu), then
//   main kernel   k_seg_direct (seg_impl 0) or k_seg_stream (1), chosen by row type and width (2): ids with <= long_thr rows
//   hot ids       k_seg_classify -> k_seg_chunk -> k_seg_long_final
// seg_fork 1 (default): the hot-id branch runs on the device's helper stream beside the main kernel
// (classify_side 1: from k_seg_classify on, so nothing precedes the main kernel on `st`).
template <int kG, int kC, typename GT, typename Epi>
int launch_direct(const void* grad, const void* grad32, int D4, const int32_t* order, const int32_t* seg,
                  const int32_t* seg_first, const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
                  const int64_t* layout, Epi epi, const DirectWs& W, cudaStream_t st) {
  const int64_t ctas = edl_cdiv(max_unique * kG, kThreads);
  EDL_REQUIRE(ctas < 0x7fffffff, EDL_ELIMIT, "segreduce: max_unique=%lld too large", (long long)max_unique);
  const int long_thr = std::max(kLongThrMin, edl_tune(EDL_TUNE_LONG_THR));
  const int batch = edl_tune(EDL_TUNE_DIRECT_BATCH);
  const int fork = edl_tune(EDL_TUNE_SEG_FORK);
  // seg_impl 2 (default) picks by row type, row width and size of the call (profiles/r2_experiments.md 7):
  // the streamed kernel for float64 / mixed rows (the owner side of the sharded step) and for wide fp32 rows
  // when there are enough ids to keep its persistent warps fed -- rows of 1 KB from 256 K ids up (dim 256, batch
  // 256K: 0.99 of the measured copy peak against 0.71), rows of 512 B from 1 M ids up -- the register path
  // otherwise (it holds too few bytes in flight per warp for wide rows, but wins on 256-byte rows and small calls).
  int impl = edl_tune(EDL_TUNE_SEG_IMPL);
  if (impl == 2) {
    if (!std::is_same<GT, float>::value) impl = 1;
    else if (D4 >= 64) impl = max_unique >= (256 << 10) ? 1 : 0;
    else if (D4 >= 32) impl = max_unique >= (1 << 20) ? 1 : 0;
    else impl = 0;
  }
  const bool cls_side = fork && edl_tune(EDL_TUNE_CLASSIFY_SIDE);
  const bool classify = fork || impl == 1;   // k_seg_direct can list the hot ids itself (unforked launch)
  std::unique_lock<std::mutex> lock(fork_mutex(), std::defer_lock);
  ForkRes* F = nullptr;
  cudaStream_t st_hot = st;
  if (fork) {
    lock.lock();
    int rc = fork_res(&F);
    if (rc) return rc;
    st_hot = F->side;
  }
  auto launch_classify = [&](cudaStream_t s) {
    k_seg_classify<<<(unsigned)edl_cdiv(max_unique, kThreads), kThreads, 0, s>>>(seg, n_unique, W.counters, W.long_list,
                                                                                W.chunk_list, long_thr);
  };
  auto launch_hot = [&]() {
    const int64_t gc = std::max<int64_t>(1, std::min<int64_t>(edl_cdiv(W.max_chunks * kG, kThreads), EDL_SM_COUNT * 8));
    k_seg_chunk<kG, kC, GT, Epi><<<(unsigned)gc, kThreads, 0, st_hot>>>((const unsigned char*)grad, (const unsigned char*)grad32, D4, order, seg,
                                                                        W.counters, W.chunk_list, W.long_list, W.partial,
                                                                        uniq, layout, epi);
    const int64_t gf = std::max<int64_t>(1, std::min<int64_t>(edl_cdiv(W.max_long / 4 * kG, kThreads), EDL_SM_COUNT * 2));
    k_seg_long_final<kG, kC, Epi><<<(unsigned)gf, kThreads, 0, st_hot>>>(D4, W.counters, W.long_list, W.partial, uniq,
                                                                         layout, epi);
  };
  if (fork) {
    if (!cls_side) launch_classify(st);
    EDL_CUDA(cudaEventRecord(F->fork, st));
    EDL_CUDA(cudaStreamWaitEvent(F->side, F->fork, 0));
    if (cls_side) launch_classify(F->side);
    launch_hot();
  } else if (classify) {
    launch_classify(st);
  }
  if (impl == 1) {
    int rc = launch_stream_main<kG, kC, GT, Epi>(grad, grad32, D4, order, seg, seg_first, n_unique, max_unique, uniq, layout,
                                                 epi, W.counters + 2, long_thr, st);
    if (rc) return rc;
  } else {
    int32_t* reg_counters = classify ? nullptr : W.counters;
#define EDL_DARGS (const unsigned char*)grad, (const unsigned char*)grad32, D4, order, seg, seg_first, n_unique, uniq, layout, epi, reg_counters, W.long_list, W.chunk_list, long_thr
    if constexpr (kC == 1) {
      if (batch == 2)
        k_seg_direct<kG, kC, GT, Epi, 8, 1, 0, kThreads, 2><<<(unsigned)edl_cdiv(ctas, 2), kThreads, 0, st>>>(EDL_DARGS);
      else
        k_seg_direct<kG, kC, GT, Epi, 8, 1><<<(unsigned)ctas, kThreads, 0, st>>>(EDL_DARGS);
    } else {
      k_seg_direct<kG, kC, GT, Epi, (kC == 2 ? 4 : 2), 1><<<(unsigned)ctas, kThreads, 0, st>>>(EDL_DARGS);
    }
#undef EDL_DARGS
  }
  if (fork) {
    EDL_CUDA(cudaEventRecord(F->join, F->side));
    EDL_CUDA(cudaStreamWaitEvent(st, F->join, 0));
  } else {
    launch_hot();
  }
  EDL_LAUNCH_CHECK("segmented reduce kernels");
  return EDL_OK;
}

// type-erased entry of this kernel family (update.cu)
inline int seg_run_direct(const SegCall& c, const DirectWs& W) {
  return seg_with_epi(c, [&](auto gt, auto epi) {
    using GT = decltype(gt);
    using Epi = decltype(epi);
    return seg_with_dim(c.dim, [&](auto G, auto C) {
      return launch_direct<decltype(G)::value, decltype(C)::value, GT, Epi>(
          c.grad, c.grad32, c.dim / 4, c.order, c.seg, c.seg_first, c.n_unique, c.max_unique, c.uniq, c.layout, epi, W, c.st);
    });
  });
}

}  // namespace
