/* edlsparse.h -- C ABI of libedlsparse.so: the B200 (sm_100a) sparse-embedding step of an
 * EasyDL ElasticTrainer parameter-server job.
 *
 * Reference interface replaced: NONE EXISTS.  The snapshot holds no PS / worker code, no
 * FFI and no operator interface.  The roles this library serves are
 * named at ref:README.md:11 (ElasticTrainer), ref:README.md:27 (recover failed parameter
 * servers and workers), ref:README.md:33-35 (scale PS / workers during training) and
 * ref:docs/design/elastic-training-operator.md:39-44,65-85 (parameter_server / worker
 * roles, replicas).  Each entry point below names the SURVEY.md section 8 row (A1..A7) and
 * the lineage function it stands in for; the semantics are written down in SPEC.md.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - the caller owns every buffer, including scratch ("workspace"): ask the matching
 *     *_workspace_bytes() function for the size.  The library allocates nothing;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - return value: 0 (EDL_OK) or a negative EDL_E* code; edl_last_error_string() gives
 *     the thread-local message.  Nothing throws, nothing aborts;
 *   - "n_dev": where a count is produced on the device (the number of unique ids) the
 *     consumers take the device pointer plus a host upper bound, so a whole step can be
 *     enqueued without a host synchronisation;
 *   - rows are fp32, row-major, `dim` a multiple of 4 (16-byte rows at least), row
 *     pointers 16-byte aligned;
 *   - "row_stride" (floats; 0 = dim): the distance between consecutive rows of a table / slot
 *     array.  A shard may keep the parameter row and the optimizer-slot rows of one id side by side
 *     in one record [p | slot0 | slot1]: slab = records, slot0 = records + dim, slot1 = records +
 *     2 dim, row_stride = (1 + n_slots) * dim.  An id's update then touches ONE DRAM page and one
 *     TLB entry instead of one per array: measured 4.39 -> 5.55 TB/s for the update's access
 *     pattern on B200 (tools/rowbw_probe), which is bound by row activations, not by bytes.
 *
 * Fused keys.  Table t has rows_t hash buckets; key(t, id) = key_base[t] + id with
 * key_base the exclusive prefix sum of rows_t.  A `layout` is a device int64 array
 *   [ T, S, shard, 0, key_base[0..T], local_base[0..T] ]      (4 + 2*(T+1) entries)
 * where shard s of table t stores local row (id div S) at slab row local_base[t] + id div S
 * and owns ids with id mod S == s (SPEC "Shard function").  layout == NULL means S == 1 and
 * slab row == key.
 */
#ifndef EDLSPARSE_H_
#define EDLSPARSE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EDL_OK 0
#define EDL_EINVAL (-1)     /* bad argument (null pointer, dim not multiple of 4, ...) */
#define EDL_ECUDA (-2)      /* a CUDA runtime call or launch failed */
#define EDL_EWORKSPACE (-3) /* workspace too small */
#define EDL_ELIMIT (-4)     /* size beyond a documented limit */

#define EDL_OPT_ADAGRAD 0
#define EDL_OPT_ADAM 1

typedef void* edl_stream_t; /* cudaStream_t */

int edl_version(void);
const char* edl_last_error_string(void);

/* ---- launch-shape knobs (no lineage counterpart: operational switches of this library) ----
 * Integer knobs that choose between kernel shapes with identical results.  Each starts at its
 * default, or at the value of its environment variable if that is set when the library first reads
 * a knob, and can be changed at any time (a launch already issued keeps what it read).
 *   long_thr      (EDL_LONG_THR, 16)      ids with more gradient rows go through the hot-id chunk kernels
 *   seg_fork      (EDL_SEG_FORK, 1)       hot-id kernels on the helper stream beside the main reduce kernel
 *   classify_side (EDL_CLASSIFY_SIDE, 1)  forked launch: the hot-id listing runs on the helper stream too
 *   seg_impl      (EDL_SEG_IMPL, 2)       main reduce kernel: 0 = k_seg_direct (register path), 1 = k_seg_stream
 *                                         (cp.async ring), 2 = k_seg_stream for float64 / mixed gradient rows and
 *                                         for fp32 rows of 128 floats and more in large calls (>= 1 M ids; >= 256 K
 *                                         ids at 256 floats), k_seg_direct otherwise (measured best for each)
 *   stream_ctas   (EDL_STREAM_CTAS, 3)    k_seg_stream: persistent CTAs per SM (capped by shared memory)
 *   stream_stages (EDL_STREAM_STAGES, 4)  k_seg_stream: ring slots in use, 2..4 (bytes in flight per warp)
 *   direct_batch  (EDL_DIRECT_BATCH, 2)   k_seg_direct: ids per lane group
 *   gather_impl   (EDL_GATHER_IMPL, 1)    gather kernels: 1 = next iteration's row index prefetched
 *   dedup_impl    (EDL_DEDUP_IMPL, 2)     edl_unique on field-major batches of at most 65536 samples: 2 = every
 *                                         field sorted by one thread-block cluster in shared memory, which also
 *                                         writes the unique step's outputs; 1 = the same sort, unique step as
 *                                         separate kernels; 0 = the global counting-sort passes
 *   dedup_shape   (EDL_DEDUP_SHAPE, 0)    that cluster: 0 = 4 CTAs x 1024 threads per field, 1 = 8 CTAs x 512 threads
 * Unknown names and out-of-range values return EDL_EINVAL. */
int edl_tuning_set(const char* name, int value);
int edl_tuning_get(const char* name, int* value_out);

/* ---- row initialisation (SPEC "Row initialisation"; lineage EmbeddingTable lazy init) ---
 * Fills slab rows [row_begin, row_begin + n_rows) of this shard's slab with the pure
 * function init(seed, t, id, c).  layout is REQUIRED here, also for S == 1: the table of a slab
 * row, and with it the id the initialiser hashes, is found through local_base. */
int edl_init_rows(float* slab, int dim, int64_t row_stride, const int64_t* layout, int64_t row_begin, int64_t n_rows,
                  uint64_t seed, edl_stream_t stream);
/* value fill for optimizer slots (Adagrad init_accum, Adam zeros) */
int edl_fill_f32(float* dst, int64_t n, float value, edl_stream_t stream);
/* the same for `dim`-wide rows that lie row_stride floats apart (a slot inside interleaved records) */
int edl_fill_rows_f32(float* dst, int dim, int64_t row_stride, int64_t n_rows, float value, edl_stream_t stream);

/* ---- A1 dedup / unique (lineage deduplicate_indexed_slices, unique half) ---------------
 * keys[j] = ids[j] + field_base[j % n_fields] (field_base NULL: keys = ids).
 * Outputs (each nullable except uniq, order, seg_offsets, n_unique):
 *   uniq[U] ascending int64, inverse[n] int32, counts[U] int32,
 *   order[n] int32 = stable argsort(keys), seg_offsets[U+1] int32 (seg_offsets[U] = n),
 *   seg_first[U] int32 (nullable) = order[seg_offsets[u]], the first batch position of each
 *   id: lets the reduce kernels skip one dependent load for ids that occur once,
 *   n_unique: device int32 scalar U.
 * Only the low key_bits bits of the keys may be non-zero (1..64).
 * field_bits (0 = off): for a fused [B, F] batch (n = B * n_fields, field_base ascending) the
 * ids of one field fit field_bits bits; the sort then runs per field on those bits only -- the
 * field is known from the position and needs no radix pass.  Same outputs either way.
 * n_dev (nullable): device int32 holding the actual n (<= n); used when the input length
 * itself was produced on the device. */
size_t edl_unique_workspace_bytes(int64_t n, int key_bits, int n_fields, int field_bits);
int edl_unique(const int64_t* ids, int64_t n, const int32_t* n_dev, const int64_t* field_base,
               int n_fields, int key_bits, int field_bits, int64_t* uniq, int32_t* inverse, int32_t* counts,
               int32_t* order, int32_t* seg_offsets, int32_t* seg_first, int32_t* n_unique,
               void* workspace, size_t workspace_bytes, edl_stream_t stream);

/* ---- A1 for callers that cannot allocate device memory -----------------------------------
 * No counterpart in the reference (nor in the ElasticDL lineage): added for this library's
 * own users.  Same inputs, outputs and results as edl_unique, but no workspace argument: the
 * scratch is a fixed EDL_UNIQUE_STATIC_SCRATCH_BYTES array inside the library image, so the
 * call reaches no cudaMalloc / cudaFree and no allocating routine of any other library.
 *
 * Who needs it: a parameter-server process whose tables were sized to fill the GPU (a raw
 * allocation may simply fail once 180 GB are spoken for); a C/C++ host that lives inside a
 * framework whose caching allocator owns all device memory and is not reachable from here;
 * and latency-critical serving threads, for which an allocation's implicit device-wide
 * synchronisation stalls every stream of the process.
 *
 * Limits and rules
 *   - n <= EDL_UNIQUE_STATIC_MAX_IDS (131072 ids, e.g. a 4096-sample x 26-field batch) and the
 *     scratch the call would need must fit the array; beyond that the call returns
 *     EDL_ELIMIT with a message and does nothing.  Use edl_unique for larger batches.
 *   - The array costs EDL_UNIQUE_STATIC_SCRATCH_BYTES of device memory on every GPU that loads
 *     the library, until the process exits.
 *   - The first call on each device looks the array's device address up (cudaGetSymbolAddress,
 *     with that device current).  Make that first call before starting any stream capture.
 *   - All calls on a device share the one array.  It still holds the previous call's data;
 *     the call never depends on it (every scratch word is written before it is read).
 *     Two calls that can overlap on the GPU (different streams, no ordering between them)
 *     race on the scratch and produce garbage: serialise them. */
#define EDL_UNIQUE_STATIC_MAX_IDS 131072
#define EDL_UNIQUE_STATIC_SCRATCH_BYTES (3 * 1024 * 1024)
int edl_unique_static(const int64_t* ids, int64_t n, const int32_t* n_dev, const int64_t* field_base,
                      int n_fields, int key_bits, int field_bits, int64_t* uniq, int32_t* inverse, int32_t* counts,
                      int32_t* order, int32_t* seg_offsets, int32_t* seg_first, int32_t* n_unique,
                      edl_stream_t stream);

/* ---- A2 hash-mod partition (lineage int_to_id / scatter_embedding_vector) --------------
 * Stable partition of keys[0..n) by (per-table id mod S'), S' = num_shards (1..255), tables
 * located through `layout` (only T and key_base are read; NULL = single table, id == key).
 * Outputs: keys_out[n] grouped by shard, offsets[S'+1] int32, perm[n] int32 (original
 * position of each output), pos[n] int32 (nullable; inverse of perm). */
size_t edl_partition_mod_workspace_bytes(int64_t n, int num_shards);
int edl_partition_mod(const int64_t* keys, int64_t n, const int32_t* n_dev, const int64_t* layout,
                      int num_shards, int64_t* keys_out, int32_t* offsets, int32_t* perm,
                      int32_t* pos, void* workspace, size_t workspace_bytes, edl_stream_t stream);

/* ---- A3 embedding gather (lineage pull_embedding_vectors / EmbeddingTable.get) ---------
 * edl_gather_keys : out[i] = slab[row(keys[i])], row() through `layout` (NULL: row = key).
 * edl_gather_ids  : out[j] = slab[ids[j] + field_base[j % n_fields]]   (S == 1 direct lookup
 *                   of a [B, F] id batch without a dedup).
 * edl_expand      : out[j] = rows[index[j]], int32 index (worker-side expand by `inverse`).
 * n_dev nullable as above. */
int edl_gather_keys(const float* slab, int dim, int64_t row_stride, const int64_t* keys, int64_t n, const int32_t* n_dev,
                    const int64_t* layout, float* out, edl_stream_t stream);
int edl_gather_ids(const float* slab, int dim, int64_t row_stride, const int64_t* ids, int64_t n,
                   const int64_t* field_base, int n_fields, float* out, edl_stream_t stream);
int edl_expand(const float* rows, int dim, const int32_t* index, int64_t n, float* out,
               edl_stream_t stream);
/* edl_expand into a window of fields of a [B, fields_total] batch: item i (sample i / fields_in_window,
 * field i % fields_in_window of the window) lands at output row
 * (i / fields_in_window) * fields_total + field0 + i % fields_in_window.  Lets several field groups,
 * looked up one after the other (pipelined exchange rounds), fill one [B * fields_total, dim] array. */
int edl_expand_window(const float* rows, int dim, const int32_t* index, int64_t n, int fields_in_window,
                      int fields_total, int field0, float* out, edl_stream_t stream);
/* index2 composition for the sharded path: out[j] = map[index[j]] (int32) */
int edl_compose_index(const int32_t* index, const int32_t* map, int64_t n, int32_t* out,
                      edl_stream_t stream);

/* ---- A4 segmented gradient reduce (lineage deduplicate_indexed_slices, sum half) -------
 * g_out[dst(u)] = sum over i in [seg_offsets[u], seg_offsets[u+1]) of grad[order[i]],
 * accumulated in float64.  dst(u) = out_pos ? out_pos[u] : u.
 * out_f64 == 0: g_out is float[.., dim], the sum rounded once to fp32.
 * out_f64 == 1: g_out is double[.., dim], the float64 partial sums themselves -- the exact
 *               wire format of the worker -> PS gradient exchange (the owner's total is then
 *               independent of how the batch was split over workers, up to float64 rounding).
 * out_f64 == 2: the exact NARROW wire (needs dst_table): an id with one gradient row is stored as that
 *               fp32 row at base32[s] (its partial sum is an fp32 value), every other id as float64 at
 *               base[s]; the receiver sums both kinds in float64 (edl_segreduce_update grad_f64 == 2),
 *               bit for bit what the all-float64 wire gives, at ~60 % of its bytes on Zipf 1.05.
 * dst_table (nullable, needs out_pos): fused reduce -> send.  Device int64 array
 *   [ G, slot_limit, offsets[0..G], base[0..G-1], base32[0..G-1] ]: output row dst(u) of group s
 *   (offsets[s] <= dst < offsets[s+1]) is stored at address base[s] + (dst - offsets[s]) rows instead of
 *   in g_out; base[s] may be another GPU's memory mapped with edl_ipc_import, so the partial sums travel
 *   over NVLink as they are made.  slot_limit > 0: rows at or beyond it within their group are dropped
 *   (the slot holds no more; the sender of the keys has flagged the overflow).  base32 is read by the
 *   mixed wire only (pass zeros otherwise).
 * n_unique_dev: device int32 U; max_unique: host upper bound on U; n_rows: number of gradient
 * rows (= length of `order`), which bounds the scratch for hot ids.
 * Streams: edl_segreduce and edl_segreduce_update run the kernels for hot ids (more than 16
 * gradient rows) on a helper stream that the library keeps per device, forked from and joined back
 * into `stream` with events, so for the caller the call is still ordered on `stream` alone and can
 * be captured into a CUDA graph through it.  The helper stream and its two events are created on
 * the first call on a device, which therefore must happen outside any stream capture.  Calls from
 * several host threads on one device are serialised over the launch sequence by a mutex.
 * Knob seg_fork = 0 (EDL_SEG_FORK=0 in the environment) keeps every launch on `stream` (no helper
 * stream is made).  The main kernel takes its shared-memory opt-in (cudaFuncSetAttribute) on the first
 * call of each shape on a device: one more reason for the first call to come before any capture. */
size_t edl_segreduce_workspace_bytes(int64_t max_unique, int64_t n_rows, int dim);
int edl_segreduce(const float* grad, int dim, const int32_t* order, const int32_t* seg_offsets,
                  const int32_t* seg_first /* nullable */, const int32_t* n_unique_dev, int64_t max_unique, int64_t n_rows,
                  const int32_t* out_pos, void* g_out, int out_f64, const int64_t* dst_table,
                  void* workspace, size_t workspace_bytes, edl_stream_t stream);

/* ---- A4 + A5 fused: reduce, then optimizer row update in place (lineage PS
 * push_gradients -> optimizer kernel) -------------------------------------------------------
 * For each unique u: g = segment sum as above, rounded once to fp32; row = row(uniq[u]) through layout.
 * Gradient rows: grad_f64 == 0: float rows in `grad`; 1: double rows in `grad` (exchanged partial sums);
 * 2: MIXED -- a row index (order / seg_first entry) with bit 30 set names a float row of `grad32`, any
 * other index a double row of `grad` (the exact narrow wire, see edl_segreduce out_f64 == 2).
 * Adagrad: acc += g*g; p -= lr*g/(sqrt(acc)+eps)                       (slot0 = acc)
 * Adam   : m = b1*m+(1-b1)*g; v = b2*v+((1-b2)*g)*g; p -= lr_t*m/(sqrt(v)+eps)
 *          (slot0 = m, slot1 = v; lr_t computed by the caller in float64, passed in `lr` or through `lr_dev`)
 * every operation rounded to fp32 separately (no FMA contraction).
 * touched (nullable): bitmap, one bit per slab row, set for every updated row (A6 export). */
typedef struct edl_opt_params {
  int kind;     /* EDL_OPT_ADAGRAD / EDL_OPT_ADAM */
  float lr;     /* Adagrad lr, or Adam lr_t */
  float eps;
  float beta1;  /* Adam only */
  float beta2;  /* Adam only */
  const float* lr_dev; /* nullable DEVICE pointer: when set, the kernels read the rate from it instead of
                        * `lr` -- a captured CUDA graph then sees the value the caller stored there before
                        * each replay (Adam's lr_t changes every step).  Zero the struct before filling it. */
} edl_opt_params;

int edl_segreduce_update(float* slab, float* slot0, float* slot1, int dim, int64_t row_stride, const int64_t* uniq,
                         const int32_t* order, const int32_t* seg_offsets,
                         const int32_t* seg_first /* nullable */, const int32_t* n_unique_dev,
                         int64_t max_unique, int64_t n_rows,
                         const void* grad, int grad_f64, const float* grad32, const int64_t* layout, const edl_opt_params* opt_host,
                         uint32_t* touched, void* workspace, size_t workspace_bytes,
                         edl_stream_t stream);

/* Self-test of the optimizer arithmetic (no lineage counterpart): the update kernels evaluate the IEEE
 * square root and division of four row elements side by side with the hardware's own fast sequences
 * and fall back to the generic routines outside their range.  This runs n pseudo-random operand
 * triples -- half raw bit patterns (zeros, denormals, infinities, NaNs), half training-like magnitudes --
 * through both forms, element-wise and composed as Adagrad / Adam steps, and ADDS the number of results
 * that differ in any bit to *mismatches_dev (device uint64, zeroed by the caller).  Expected: 0. */
int edl_selftest_rounding(int64_t n, uint64_t seed, uint64_t* mismatches_dev, edl_stream_t stream);

/* ---- A6 elastic re-shard (lineage CheckpointSaver / restore_params_from_checkpoint) ----
 * edl_export_count : n_out (device int32) = number of set bits in touched[0..n_rows).
 * edl_export_shard : compacts the touched rows in ascending slab-row order into
 *                    keys_out[n] (fused keys), rows_out / slot0_out / slot1_out [n, dim].
 *                    capacity = size of the output buffers in rows.
 * edl_import_shard : for each record i with (id mod S == shard) under `layout`: writes
 *                    rows/slots into the slab and sets its touched bit; others are skipped
 *                    (restore reads ALL old shard files and keeps its own ids). */
size_t edl_export_workspace_bytes(int64_t n_rows);
int edl_export_count(const uint32_t* touched, int64_t n_rows, int32_t* n_out, void* workspace,
                     size_t workspace_bytes, edl_stream_t stream);
int edl_export_shard(const float* slab, const float* slot0, const float* slot1, int dim, int64_t row_stride,
                     const uint32_t* touched, int64_t n_rows, const int64_t* layout,
                     int64_t capacity, int64_t* keys_out, float* rows_out, float* slot0_out,
                     float* slot1_out, int32_t* n_out, void* workspace, size_t workspace_bytes,
                     edl_stream_t stream);
int edl_import_shard(float* slab, float* slot0, float* slot1, int dim, int64_t row_stride, uint32_t* touched,
                     const int64_t* layout, const int64_t* keys, const float* rows,
                     const float* slot0_in, const float* slot1_in, int64_t n,
                     edl_stream_t stream);

/* ---- A7 exchange helpers: permute rows for the all-to-all-v send buffers --------------
 * out[i] = rows[perm[i]] is edl_expand.  Scatter form: out[pos[i]] = rows[i]. */
int edl_scatter_rows(const float* rows, int dim, const int32_t* pos, int64_t n,
                     const int32_t* n_dev, float* out, int64_t out_row_stride /* floats; 0 = dim */, edl_stream_t stream);

/* ---- lazily materialised rows (lineage EmbeddingTable.get lazy initialisation) --
 * For tables that do not fit densely (configs[2]: 1.6 TB with Adam state).  A shard stores only
 * rows that received an update, in a slab of max_rows + 1 rows (the last one is a spare that
 * absorbs writes after an overflow), behind an open-addressing hash index
 * fused key -> slab row.  The caller owns both: `index` is edl_lazy_index_bytes(index_capacity)
 * bytes (index_capacity a power of two >= 2 x max_rows), cleared with edl_lazy_clear.
 *   edl_lazy_gather : out[i] = stored row of keys[i], or the initial value init(seed, t, id)
 *                     computed on the fly when the row was never updated (nothing is inserted).
 *   edl_lazy_insert : for UNIQUE keys (the output of edl_unique): slots_out[i] = slab row of
 *                     keys[i]; unseen keys claim the next free row, which is initialised
 *                     (row = init, slot0 / slot1 = the given constants).  Feed slots_out as
 *                     `uniq` with layout == NULL to edl_segreduce_update.
 *   edl_lazy_status_ptr : device int32[2] = {rows materialised, overflow flag}.
 *   edl_lazy_dump   : keys_out[s], slots_out[s] for every materialised row s (slot order). */
size_t edl_lazy_index_bytes(int64_t index_capacity);
int edl_lazy_clear(void* index, int64_t index_capacity, edl_stream_t stream);
int edl_lazy_gather(const void* index, int64_t index_capacity, const float* slab, int dim, int64_t row_stride,
                    const int64_t* keys, int64_t n, const int32_t* n_dev, const int64_t* layout,
                    uint64_t seed, float* out, edl_stream_t stream);
int edl_lazy_insert(void* index, int64_t index_capacity, float* slab, float* slot0, float* slot1,
                    int dim, int64_t row_stride, int64_t max_rows, const int64_t* keys, int64_t n, const int32_t* n_dev,
                    const int64_t* layout, uint64_t seed, float slot0_init, float slot1_init,
                    int64_t* slots_out, edl_stream_t stream);
/* stamps[slots[i]] = value for i < *n_dev: marks the rows of one update (slots = edl_lazy_insert's output)
 * with the step number, for an eviction policy that drops the rows not updated for longest. */
int edl_lazy_stamp(int32_t* stamps, const int64_t* slots, int64_t n, const int32_t* n_dev, int32_t value,
                   edl_stream_t stream);
const int32_t* edl_lazy_status_ptr(const void* index, int64_t index_capacity);
int edl_lazy_dump(const void* index, int64_t index_capacity, int64_t* keys_out, int64_t* slots_out,
                  edl_stream_t stream);

/* ---- A7 over NVLink peer memory (replaces the worker <-> PS RPC of the lineage) ----------
 * Exchange buffers that the other ranks of the box write into directly.  Set-up: every rank
 * allocates its receive buffer (edl_peer_alloc: the one place this library allocates, because a
 * CUDA IPC handle exists only for a whole cudaMalloc allocation), exports a 64-byte handle
 * (edl_ipc_export), ships it to its peers by any host channel, and maps theirs
 * (edl_ipc_import).  Per step: edl_peer_copy into the mapped pointers (NVLink), or kernels of
 * this library given a mapped pointer as their output. */
int edl_peer_alloc(void** ptr, size_t bytes);
int edl_peer_free(void* ptr);
int edl_ipc_export(void* ptr, unsigned char* handle64);
int edl_ipc_import(const unsigned char* handle64, void** ptr);
int edl_ipc_close(void* ptr);
int edl_peer_copy(void* dst, const void* src, size_t bytes, edl_stream_t stream);
/* One kernel per rank that closes an exchange round: publishes `value` into slot `me` of every
 * peer's flag array (peer_flag_ptrs_dev: device array of `world` mapped pointers), then waits
 * until all `world` slots of own_flags hold >= value.  own_flags has world + 1 entries; entry
 * [world] is set to -1 if the wait timed out.  One rank per GPU. */
int edl_peer_signal_wait(const int64_t* peer_flag_ptrs_dev, int64_t* own_flags, int me, int world,
                         int64_t value, edl_stream_t stream);
/* The same with the round number kept on the device: the kernel increments *counter_dev and
 * publishes the new value, so a CUDA graph that contains it signals a fresh round on every replay. */
int edl_peer_signal_wait_counter(const int64_t* peer_flag_ptrs_dev, int64_t* own_flags, int me, int world,
                                 int64_t* counter_dev, edl_stream_t stream);

/* ---- A7 rounds over fixed per-pair slots: no size ever reaches the host -------------------
 * (replaces the worker <-> PS RPC of the lineage, like the section above; no reference file.)
 * A receive buffer is `num_slots` (= world size) slots of `slot_stride` rows; slot s is written only
 * by rank s, so a sender needs only its own partition offsets, which edl_partition_mod leaves on
 * the device.  Row counts travel with the data.  Every launch below has a fixed grid, so the whole
 * sharded step can be captured into a CUDA graph.
 *
 * edl_slot_send    : rows[p] for p in [offsets[d], offsets[d+1]) -> dst_bases[d] + (p - offsets[d]) rows
 *                    (dst_bases_dev: device array of num_slots addresses, usually this rank's slot in
 *                    every peer's buffer); count_ptrs_dev (nullable): device array of num_slots
 *                    addresses of int32 cells that receive the group sizes.  Groups longer than
 *                    slot_limit are cut there and *err_dev |= 1 (check it after the step).
 *                    n_max: host bound on offsets[num_slots] (sizes the grid only).
 *                    flag_seg_offsets / flag_perm (nullable; 8-byte keys only): row p is unique id
 *                    u = flag_perm[p] of an edl_unique result with segment offsets flag_seg_offsets; keys
 *                    of ids with exactly ONE gradient row are sent with bit 62 set (the narrow wire: the
 *                    owner will find that id's partial sum as an fp32 row).
 * edl_slot_positions: out[u] = d * slot_stride + min(pos[u] - offsets[d], slot_stride - 1), the
 *                    slot-space row of partition position pos[u] (receiver side of the rows round).
 * edl_slot_compact : owner side: slots of keys with counts[s] valid rows -> contiguous keys_out in
 *                    slot (= peer rank) order with bit 62 cleared, map_out[i] = slot-space row of
 *                    keys_out[i], with bit 30 set where the key carried bit 62, and *n_out.
 * edl_remap_index  : out[i] = map[index[i]] for i < *n_dev (edl_compose_index with a device count):
 *                    turns the `order` / `seg_first` of edl_unique(keys_out) into slot-space rows.
 * edl_slot_merge_unique: the three calls above in one (edl_slot_compact -> edl_unique -> edl_remap_index x 2) for
 *                    slots whose keys are strictly increasing (what edl_partition_mod + edl_slot_send deliver from
 *                    an edl_unique result): the runs are MERGED (S - 1 binary searches per key) instead of sorted.
 *                    uniq / seg_offsets / n_unique as edl_unique of the concatenated keys would give them; order and
 *                    seg_first already as slot-space rows (bit 30 = the key carried bit 62); *n_out_dev = keys
 *                    received.  Arrays of num_slots * slot_limit entries (seg_offsets one more).  Keys out of
 *                    order inside a slot give an unspecified result: the caller guarantees the precondition. */
int edl_slot_send(const void* rows, int row_bytes, const int32_t* offsets_dev, int num_slots,
                  const int64_t* dst_bases_dev, int64_t slot_limit, const int64_t* count_ptrs_dev,
                  int32_t* err_dev, int64_t n_max, const int32_t* flag_seg_offsets, const int32_t* flag_perm,
                  edl_stream_t stream);
int edl_slot_positions(const int32_t* pos, int64_t n_max, const int32_t* n_dev, const int32_t* offsets_dev,
                       int num_slots, int64_t slot_stride, int32_t* out, edl_stream_t stream);
size_t edl_slot_merge_unique_workspace_bytes(int num_slots, int64_t slot_limit);
int edl_slot_merge_unique(const int64_t* keys_slots, const int32_t* counts_dev, int num_slots, int64_t slot_stride,
                          int64_t slot_limit, int64_t* uniq, int32_t* order, int32_t* seg_offsets, int32_t* seg_first,
                          int32_t* n_unique, int32_t* n_out_dev, void* workspace, size_t workspace_bytes,
                          edl_stream_t stream);
int edl_slot_compact(const int64_t* keys_slots, const int32_t* counts_dev, int num_slots, int64_t slot_stride,
                     int64_t slot_limit, int64_t* keys_out, int32_t* map_out, int32_t* n_out_dev,
                     edl_stream_t stream);
int edl_remap_index(const int32_t* index, const int32_t* map, int64_t n_max, const int32_t* n_dev,
                    int32_t* out, edl_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* EDLSPARSE_H_ */
