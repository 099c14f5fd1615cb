// A4 segmented sparse-gradient reduce and A5 Adagrad / Adam row update (fused): the C entry
// points and the choice between the two kernel families.
//
// SPEC: g_uniq[u] = sum of the gradient rows that share id u,
// accumulated in float64 and rounded once to fp32; then the optimizer update on the touched
// row with every fp32 operation rounded separately (explicit __f*_rn, so nvcc cannot contract
// to FMA and the result matches the NumPy oracle op for op).
//
//   seg_stream.cuh  main kernel (default): persistent warps, rows streamed through a private
//                   shared-memory ring with cp.async
//   seg_direct.cuh  the register-path main kernel (knob seg_impl = 0), the hot-id kernels and the launch
#include "seg_common.cuh"
#include "seg_direct.cuh"

namespace {

struct Ws { DirectWs direct; size_t zero_bytes; size_t bytes; };

// [counters] cleared per call, then the hot-id lists, then the float64 partial rows.  Partial rows exist
// only for ids that span more than one chunk (more than kChunk gradient rows): at most
// n / kChunk + n / (kChunk + 1) chunks belong to such ids.
Ws carve_ws(void* ws, int64_t n_rows, int dim) {
  Ws w; char* p = (char*)ws;
  size_t off = 0;
  w.direct.counters = (int32_t*)(p + off); off += 256;
  w.zero_bytes = off;
  const int64_t n = n_rows > 0 ? n_rows : 0;
  w.direct.max_long = n / (kLongThrMin + 1) + 1;
  w.direct.max_chunks = n / kChunk + w.direct.max_long + 1;
  const int64_t max_partial = n / kChunk + n / (kChunk + 1) + 2;
  w.direct.long_list = (int32_t*)(p + off); off = edl_align_up(off + (size_t)w.direct.max_long * 4 * 4, 256);
  w.direct.chunk_list = (int32_t*)(p + off); off = edl_align_up(off + (size_t)w.direct.max_chunks * 4 * 4, 256);
  w.direct.partial = (double*)(p + off); off = edl_align_up(off + (size_t)max_partial * dim * 8, 256);
  w.bytes = off;
  return w;
}

int run_call(const SegCall& c, const Ws& W) {
  EDL_CUDA(cudaMemsetAsync(W.direct.counters, 0, W.zero_bytes, c.st));
  return seg_run_direct(c, W.direct);
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
  EDL_REQUIRE(out_f64 >= 0 && out_f64 <= 2, EDL_EINVAL, "edl_segreduce: out_f64=%d not 0 (fp32), 1 (float64) or 2 (mixed)", out_f64);
  EDL_REQUIRE(out_f64 != 2 || dst_table, EDL_EINVAL, "edl_segreduce: the mixed wire (out_f64 = 2) needs a dst_table");
  SegCall c{};
  c.grad = grad; c.dim = dim; c.grad_f64 = 0; c.order = order; c.seg = seg_offsets; c.seg_first = seg_first;
  c.n_unique = n_unique_dev; c.max_unique = max_unique; c.uniq = nullptr; c.layout = nullptr;
  c.epi_kind = out_f64 == 2 ? 4 : (out_f64 ? 1 : 0); c.g_out = g_out; c.out_pos = out_pos; c.dst_table = dst_table;
  c.st = (cudaStream_t)stream;
  return run_call(c, W);
}

extern "C" int edl_segreduce_update(float* slab, float* slot0, float* slot1, int dim, int64_t row_stride,
                                    const int64_t* uniq, const int32_t* order,
                                    const int32_t* seg_offsets, const int32_t* seg_first,
                                    const int32_t* n_unique_dev,
                                    int64_t max_unique, int64_t n_rows, const void* grad, int grad_f64,
                                    const float* grad32, const int64_t* layout, const edl_opt_params* opt,
                                    uint32_t* touched, void* workspace, size_t workspace_bytes,
                                    edl_stream_t stream) {
  if (max_unique == 0) return EDL_OK;
  Ws W;
  int rc = check_common("edl_segreduce_update", grad, dim, order, seg_offsets, n_unique_dev,
                        max_unique, n_rows, workspace, workspace_bytes, &W);
  if (rc) return rc;
  EDL_REQUIRE(opt, EDL_EINVAL, "edl_segreduce_update: opt is null");
  int RS4 = 0;
  { int rc0 = edl_row_stride4("edl_segreduce_update", dim, row_stride, &RS4); if (rc0) return rc0; }
  EDL_REQUIRE(slab && slot0 && uniq, EDL_EINVAL, "edl_segreduce_update: null slab / slot0 / uniq");
  EDL_REQUIRE((((uintptr_t)slab | (uintptr_t)slot0 | (uintptr_t)slot1) & 15) == 0, EDL_EINVAL,
              "edl_segreduce_update: rows must be 16-byte aligned");
  OptDev o;
  o.lr_dev = opt->lr_dev;
  o.lr = opt->lr; o.eps = opt->eps; o.b1 = opt->beta1; o.b2 = opt->beta2;
  o.omb1 = 1.0f - opt->beta1;  // fp32 subtraction, as the oracle does
  o.omb2 = 1.0f - opt->beta2;
  EDL_REQUIRE(opt->kind == EDL_OPT_ADAGRAD || opt->kind == EDL_OPT_ADAM, EDL_EINVAL,
              "edl_segreduce_update: unknown optimizer kind %d", opt->kind);
  EDL_REQUIRE(opt->kind != EDL_OPT_ADAM || slot1, EDL_EINVAL, "edl_segreduce_update: Adam needs slot1 (v)");
  SegCall c{};
  EDL_REQUIRE(grad_f64 >= 0 && grad_f64 <= 2, EDL_EINVAL, "edl_segreduce_update: grad_f64=%d not 0, 1 or 2", grad_f64);
  EDL_REQUIRE(grad_f64 != 2 || (grad32 && (((uintptr_t)grad32) & 15) == 0), EDL_EINVAL,
              "edl_segreduce_update: mixed gradient rows (grad_f64 = 2) need grad32");
  c.grad = grad; c.grad32 = grad32; c.dim = dim; c.grad_f64 = grad_f64; c.order = order; c.seg = seg_offsets; c.seg_first = seg_first;
  c.n_unique = n_unique_dev; c.max_unique = max_unique; c.uniq = uniq; c.layout = layout;
  c.epi_kind = opt->kind == EDL_OPT_ADAGRAD ? 2 : 3;
  c.slab = slab; c.slot0 = slot0; c.slot1 = slot1; c.touched = touched; c.opt = o; c.rs4 = RS4;
  c.st = (cudaStream_t)stream;
  return run_call(c, W);
}

// ---- self-test of the four-wide IEEE sqrt / divide sequences (seg_common.cuh) --------------------------
namespace {
__device__ __forceinline__ unsigned long long sm64(unsigned long long x) {
  unsigned long long z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// mode 0: raw random bit patterns (every class of float: zeros, denormals, infinities, NaNs, huge, tiny)
// mode 1: positive values with exponents spread over 2^-40 .. 2^40 (what optimizer state looks like)
__device__ __forceinline__ float rnd_float(unsigned long long h, int mode) {
  if (mode == 0) return __uint_as_float((unsigned)h);
  const unsigned e = 87u + (unsigned)((h >> 40) % 81ull);
  return __uint_as_float((e << 23) | ((unsigned)h & 0x7fffffu) | ((unsigned)(h >> 63) << 31));
}
__global__ void __launch_bounds__(256)
k_selftest_rounding(long long n, unsigned long long seed, unsigned long long* __restrict__ bad) {
  unsigned long long mine = 0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const int mode = (int)(i & 1);
    const unsigned long long h0 = sm64(seed ^ (unsigned long long)i), h1 = sm64(h0), h2 = sm64(h1);
    const float a = fabsf(rnd_float(h0, mode)), x = rnd_float(h1, mode), d = rnd_float(h2, mode);
    const float s_ref = __fsqrt_rn(a);
    const float s_new = sqrt_fast_ok(a) ? sqrt_fast(a) : __fsqrt_rn(a);
    const float q_ref = __fdiv_rn(x, d);
    const float q_new = div_fast_ok(x, d) ? div_fast(x, d) : __fdiv_rn(x, d);
    mine += (__float_as_uint(s_ref) != __float_as_uint(s_new)) + (__float_as_uint(q_ref) != __float_as_uint(q_new));
    // and the composed forms the optimizers use: every branch of adagrad4 against four adagrad1
    OptDev o{0.01f, 1e-7f, 0.9f, 0.999f, 0.1f, 0.001f, nullptr};
    float4 g = make_float4(x, d, __uint_as_float((unsigned)(h2 >> 32)), 0.0f);
    if (mode == 1) g = make_float4(x * 1e-3f, d * 1e-3f, -x, 0.0f);
    float4 p = make_float4(0.01f, -0.02f, x, -0.0f), acc = make_float4(0.1f, a, a + 0.1f, 0.1f);
    float4 p1 = p, acc1 = acc;
    adagrad4(g, p1, acc1, o);
    adagrad1(g.x, p.x, acc.x, o); adagrad1(g.y, p.y, acc.y, o); adagrad1(g.z, p.z, acc.z, o); adagrad1(g.w, p.w, acc.w, o);
    mine += (__float_as_uint(p.x) != __float_as_uint(p1.x)) + (__float_as_uint(p.y) != __float_as_uint(p1.y)) +
            (__float_as_uint(p.z) != __float_as_uint(p1.z)) + (__float_as_uint(p.w) != __float_as_uint(p1.w)) +
            (__float_as_uint(acc.x) != __float_as_uint(acc1.x)) + (__float_as_uint(acc.y) != __float_as_uint(acc1.y)) +
            (__float_as_uint(acc.z) != __float_as_uint(acc1.z)) + (__float_as_uint(acc.w) != __float_as_uint(acc1.w));
    float4 pm = make_float4(0.01f, -0.02f, x, -0.0f), m = make_float4(0.0f, x * 1e-2f, -d, 0.0f), v = make_float4(0.0f, a, a * 1e-6f, 1e-30f);
    float4 pm1 = pm, m1 = m, v1 = v;
    adam4(g, pm1, m1, v1, o);
    adam1(g.x, pm.x, m.x, v.x, o); adam1(g.y, pm.y, m.y, v.y, o); adam1(g.z, pm.z, m.z, v.z, o); adam1(g.w, pm.w, m.w, v.w, o);
    mine += (__float_as_uint(pm.x) != __float_as_uint(pm1.x)) + (__float_as_uint(pm.y) != __float_as_uint(pm1.y)) +
            (__float_as_uint(pm.z) != __float_as_uint(pm1.z)) + (__float_as_uint(pm.w) != __float_as_uint(pm1.w)) +
            (__float_as_uint(m.x) != __float_as_uint(m1.x)) + (__float_as_uint(m.y) != __float_as_uint(m1.y)) +
            (__float_as_uint(m.z) != __float_as_uint(m1.z)) + (__float_as_uint(m.w) != __float_as_uint(m1.w)) +
            (__float_as_uint(v.x) != __float_as_uint(v1.x)) + (__float_as_uint(v.y) != __float_as_uint(v1.y)) +
            (__float_as_uint(v.z) != __float_as_uint(v1.z)) + (__float_as_uint(v.w) != __float_as_uint(v1.w));
  }
  if (mine) atomicAdd(bad, mine);
}
}  // namespace

extern "C" int edl_selftest_rounding(int64_t n, uint64_t seed, uint64_t* mismatches_dev, edl_stream_t stream) {
  EDL_REQUIRE(n >= 0 && mismatches_dev, EDL_EINVAL, "edl_selftest_rounding: bad argument");
  if (n == 0) return EDL_OK;
  const int g = (int)std::min<int64_t>(edl_cdiv(n, 256), EDL_SM_COUNT * 16);
  k_selftest_rounding<<<g, 256, 0, (cudaStream_t)stream>>>((long long)n, (unsigned long long)seed,
                                                           (unsigned long long*)mismatches_dev);
  EDL_LAUNCH_CHECK("k_selftest_rounding");
  return EDL_OK;
}
