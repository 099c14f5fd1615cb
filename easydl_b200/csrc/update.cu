// A4 segmented sparse-gradient reduce and A5 Adagrad / Adam row update (fused): the C entry
// points and the choice between the two kernel families.
//
// SPEC: g_uniq[u] = sum of the gradient rows that share id u,
// accumulated in float64 and rounded once to fp32; then the optimizer update on the touched
// row with every fp32 operation rounded separately (explicit __f*_rn, so nvcc cannot contract
// to FMA and the result matches the NumPy oracle op for op).
//
//   seg_direct.cuh  register path: one lane group per id, hot ids cut into chunks (default)
//   seg_tile.cuh    shared-memory staged path: persistent CTAs over the sorted positions,
//                   asynchronous row copies into a 3-stage ring (EDL_SEG_IMPL=tile)
#include "seg_common.cuh"
#include "seg_direct.cuh"
#include "seg_tile.cuh"

namespace {

struct Ws { SegWs tile; DirectWs direct; size_t zero_bytes; size_t bytes; };

// [flags | ticket | direct counters] cleared per call, then tile carries, then direct lists
Ws carve_ws(void* ws, int64_t n_rows, int dim) {
  Ws w; char* p = (char*)ws;
  size_t off = 0;
  w.tile.flags = (int32_t*)(p + off); off += (size_t)kMaxGrid * 4 * sizeof(int32_t);
  w.tile.ticket = (unsigned*)(p + off); off += 256;
  w.direct.counters = (int32_t*)(p + off); off += 256;
  w.zero_bytes = off;
  w.tile.zero_bytes = off;
  w.tile.carry_h = (double*)(p + off); off += (size_t)kMaxGrid * dim * sizeof(double);
  w.tile.carry_t = (double*)(p + off); off += (size_t)kMaxGrid * dim * sizeof(double);
  const int64_t n = n_rows > 0 ? n_rows : 0;
  w.direct.max_long = n / (kLongThrMin + 1) + 1;
  w.direct.max_chunks = n / kChunk + w.direct.max_long + 1;
  w.direct.long_list = (int32_t*)(p + off); off = edl_align_up(off + (size_t)w.direct.max_long * 4 * 4, 256);
  w.direct.chunk_list = (int32_t*)(p + off); off = edl_align_up(off + (size_t)w.direct.max_chunks * 4 * 4, 256);
  w.direct.partial = (double*)(p + off); off = edl_align_up(off + (size_t)w.direct.max_chunks * dim * 8, 256);
  w.tile.bytes = off;
  w.bytes = off;
  return w;
}

bool use_tile_impl() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("EDL_SEG_IMPL");
    v = (e && e[0] == 't') ? 1 : 0;
  }
  return v == 1;
}

template <typename GT, typename Epi>
int dispatch_dim(const void* grad, int dim, const int32_t* order, const int32_t* seg,
                 const int32_t* seg_first, const int32_t* n_unique, int64_t max_unique, const int64_t* uniq,
                 const int64_t* layout, Epi epi, const Ws& W, cudaStream_t st) {
  const int D4 = dim / 4;
  EDL_CUDA(cudaMemsetAsync(W.tile.flags, 0, W.zero_bytes, st));
  const bool tile = use_tile_impl();
#define EDL_GO(G, C)                                                                                          \
  return tile ? launch_tile<G, C, GT, Epi>(grad, D4, order, seg, n_unique, max_unique, uniq, layout, epi, W.tile, st) \
              : launch_direct<G, C, GT, Epi>(grad, D4, order, seg, seg_first, n_unique, max_unique, uniq, layout, epi, W.direct, st)
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
                 const int32_t* seg, const int32_t* n_unique, int64_t max_unique, int64_t n_rows,
                 void* ws, size_t ws_bytes, Ws* W) {
  EDL_REQUIRE(dim > 0 && dim % 4 == 0, EDL_EINVAL, "%s: dim=%d must be a positive multiple of 4", who, dim);
  EDL_REQUIRE(max_unique >= 0, EDL_EINVAL, "%s: max_unique negative", who);
  EDL_REQUIRE(grad && order && seg && n_unique, EDL_EINVAL, "%s: null pointer", who);
  EDL_REQUIRE((((uintptr_t)grad) & 15) == 0, EDL_EINVAL, "%s: grad must be 16-byte aligned", who);
  EDL_REQUIRE(n_rows >= 0 && n_rows < (int64_t)0x7fff0000, EDL_EINVAL, "%s: n_rows=%lld out of range", who, (long long)n_rows);
  *W = carve_ws(ws, n_rows, dim);
  EDL_REQUIRE(ws && ws_bytes >= W->bytes, EDL_EWORKSPACE, "%s: workspace %zu < required %zu", who,
              ws_bytes, W->bytes);
  return EDL_OK;
}

}  // namespace

extern "C" size_t edl_segreduce_workspace_bytes(int64_t max_unique, int64_t n_rows, int dim) {
  (void)max_unique;
  return carve_ws(nullptr, n_rows, dim > 0 ? dim : 4).bytes;
}

extern "C" int edl_segreduce(const float* grad, int dim, const int32_t* order,
                             const int32_t* seg_offsets, const int32_t* seg_first,
                             const int32_t* n_unique_dev,
                             int64_t max_unique, int64_t n_rows, const int32_t* out_pos, void* g_out, int out_f64,
                             const int64_t* dst_table, void* workspace, size_t workspace_bytes,
                             edl_stream_t stream) {
  if (max_unique == 0) return EDL_OK;
  Ws W;
  int rc = check_common("edl_segreduce", grad, dim, order, seg_offsets, n_unique_dev, max_unique, n_rows,
                        workspace, workspace_bytes, &W);
  if (rc) return rc;
  EDL_REQUIRE(dst_table || (g_out && (((uintptr_t)g_out) & 15) == 0), EDL_EINVAL,
              "edl_segreduce: g_out null or misaligned (and no dst_table)");
  EDL_REQUIRE(!dst_table || out_pos, EDL_EINVAL, "edl_segreduce: dst_table needs out_pos");
  cudaStream_t st = (cudaStream_t)stream;
  if (out_f64) {
    EpiWrite64 epi{(double*)g_out, out_pos, dst_table};
    return dispatch_dim<float, EpiWrite64>(grad, dim, order, seg_offsets, seg_first, n_unique_dev, max_unique, nullptr,
                                           nullptr, epi, W, st);
  }
  EpiWrite epi{(float4*)g_out, out_pos, dst_table};
  return dispatch_dim<float, EpiWrite>(grad, dim, order, seg_offsets, seg_first, n_unique_dev, max_unique, nullptr,
                                       nullptr, epi, W, st);
}

extern "C" int edl_segreduce_update(float* slab, float* slot0, float* slot1, int dim,
                                    const int64_t* uniq, const int32_t* order,
                                    const int32_t* seg_offsets, const int32_t* seg_first,
                                    const int32_t* n_unique_dev,
                                    int64_t max_unique, int64_t n_rows, const void* grad, int grad_f64,
                                    const int64_t* layout, const edl_opt_params* opt,
                                    uint32_t* touched, void* workspace, size_t workspace_bytes,
                                    edl_stream_t stream) {
  if (max_unique == 0) return EDL_OK;
  Ws W;
  int rc = check_common("edl_segreduce_update", grad, dim, order, seg_offsets, n_unique_dev,
                        max_unique, n_rows, workspace, workspace_bytes, &W);
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
    return grad_f64 ? dispatch_dim<double, EpiAdagrad>(grad, dim, order, seg_offsets, seg_first, n_unique_dev, max_unique, uniq, layout, epi, W, st)
                    : dispatch_dim<float, EpiAdagrad>(grad, dim, order, seg_offsets, seg_first, n_unique_dev, max_unique, uniq, layout, epi, W, st);
  }
  if (opt->kind == EDL_OPT_ADAM) {
    EDL_REQUIRE(slot1, EDL_EINVAL, "edl_segreduce_update: Adam needs slot1 (v)");
    EpiAdam epi{(float4*)slab, (float4*)slot0, (float4*)slot1, touched, o};
    return grad_f64 ? dispatch_dim<double, EpiAdam>(grad, dim, order, seg_offsets, seg_first, n_unique_dev, max_unique, uniq, layout, epi, W, st)
                    : dispatch_dim<float, EpiAdam>(grad, dim, order, seg_offsets, seg_first, n_unique_dev, max_unique, uniq, layout, epi, W, st);
  }
  edl_set_error("edl_segreduce_update: unknown optimizer kind %d", opt->kind);
  return EDL_EINVAL;
}
