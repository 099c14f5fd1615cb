/* edl_oracle_c.c -- second CPU restatement of SPEC.md, in plain C.  TEST INFRASTRUCTURE.
 *
 * PARITY UNPINNED, like oracle/edl_oracle.py: the reference snapshot holds no implementation, tests or
 * golden vectors for this path (ref:README.md:11 is the only mention of the
 * ElasticTrainer), so there is no reference file for any function below to follow -- each one cites the
 * SPEC.md clause it restates.  What this file adds is independence: it shares no code with the NumPy
 * oracle (its own radix sort instead of np.unique / argsort, its own hash-table row store instead of
 * searchsorted over sorted arrays, scalar loops instead of ufuncs), and tests/test_oracle_c.py requires the
 * two to agree bit for bit on every output, including multi-step Adagrad / Adam trajectories.  A
 * disagreement means one of the restatements misreads the SPEC.
 *
 * It is also the compiled, multi-threaded form of the CPU parameter-server path (one table per thread, a
 * hash map id -> row per table, as a PS process keeps its embedding tables), which bench.py reports as
 * `cpu_baseline_compiled` beside the NumPy figure.
 *
 * Only tests/, __graft_entry__ (build) and bench.py's CPU legs may load this library.  The product
 * (easydl_b200/) never does.  Build: make -C oracle   (gcc -O2 -ffp-contract=off: no FMA contraction, so
 * every fp32 operation is rounded on its own as SPEC 'A5 operation order' demands).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define EDLO_ADAGRAD 0
#define EDLO_ADAM 1

/* The inner loops are compiled twice (SSE2 and AVX2, picked at load time by the CPU): same IEEE operations,
 * same order per element, no FMA in either (-ffp-contract=off, and the avx2 target does not enable fma),
 * so the results do not depend on which clone runs. */
#if defined(__x86_64__) && defined(__GNUC__)
#define EDLO_CLONES __attribute__((target_clones("avx2", "default")))
#else
#define EDLO_CLONES
#endif

/* gradient rows of a segment lie anywhere in the batch: ask for the row a few iterations ahead */
static inline void prefetch_row(const float* row, int dim) {
  for (int c = 0; c < dim; c += 16) __builtin_prefetch(row + c, 0, 0);
}

EDLO_CLONES static void acc_add(double* restrict acc, const float* restrict g, int dim) {
  for (int c = 0; c < dim; ++c) acc[c] += (double)g[c];
}
EDLO_CLONES static void acc_round(const double* restrict acc, float* restrict out, int dim) {
  for (int c = 0; c < dim; ++c) out[c] = (float)acc[c];
}

/* ------------------------------------------------------------------ SPEC 'Row initialisation' */
uint64_t edlo_splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

static void init_row(uint64_t seed, int t, int64_t id, int dim, float* out) {
  const uint64_t salt = seed ^ ((uint64_t)t << 56);
  for (int c = 0; c < dim; ++c) {
    const uint64_t h = edlo_splitmix64(salt ^ ((uint64_t)id * (uint64_t)dim + (uint64_t)c));
    const float x = (float)(h >> 40) * 0x1p-24f; /* 24 bits: exact */
    const float w = x * 0.1f;                     /* two separately rounded operations */
    out[c] = w - 0.05f;
  }
}

void edlo_init_rows(uint64_t seed, int t, const int64_t* ids, int64_t n, int dim, float* out) {
  for (int64_t i = 0; i < n; ++i) init_row(seed, t, ids[i], dim, out + i * dim);
}

/* ------------------------------------------------------------------ SPEC 'A1 dedup'
 * Stable LSD radix sort of (key, position) pairs, 8-bit digits, passes whose digit is the same for
 * every key skipped; then one scan for uniq / counts / seg_offsets / inverse.
 * -> U.  order[n] = stable argsort, seg[U + 1], inverse[n], counts[U], uniq[U] (capacity n each). */
int64_t edlo_dedup(const int64_t* keys, int64_t n, int64_t* uniq, int32_t* inverse, int32_t* counts,
                   int32_t* order, int32_t* seg) {
  seg[0] = 0;
  if (n <= 0) return 0;
  uint64_t* ka = (uint64_t*)malloc((size_t)n * 8);
  uint64_t* kb = (uint64_t*)malloc((size_t)n * 8);
  int32_t* pa = (int32_t*)malloc((size_t)n * 4);
  int32_t* pb = (int32_t*)malloc((size_t)n * 4);
  uint64_t any = 0, all = ~0ull;
  for (int64_t i = 0; i < n; ++i) {
    ka[i] = (uint64_t)keys[i] ^ 0x8000000000000000ull; /* order-preserving map of int64 onto uint64 */
    pa[i] = (int32_t)i;
    any |= ka[i];
    all &= ka[i];
  }
  for (int shift = 0; shift < 64; shift += 8) {
    if ((((any ^ all) >> shift) & 0xff) == 0) continue; /* this digit is equal everywhere */
    int64_t hist[257];
    memset(hist, 0, sizeof hist);
    for (int64_t i = 0; i < n; ++i) hist[((ka[i] >> shift) & 0xff) + 1]++;
    for (int d = 0; d < 256; ++d) hist[d + 1] += hist[d];
    for (int64_t i = 0; i < n; ++i) {
      const int64_t dst = hist[(ka[i] >> shift) & 0xff]++;
      kb[dst] = ka[i];
      pb[dst] = pa[i];
    }
    uint64_t* tk = ka; ka = kb; kb = tk;
    int32_t* tp = pa; pa = pb; pb = tp;
  }
  int64_t U = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (i == 0 || ka[i] != ka[i - 1]) {
      uniq[U] = (int64_t)(ka[i] ^ 0x8000000000000000ull);
      seg[U] = (int32_t)i;
      ++U;
    }
    order[i] = pa[i];
    inverse[pa[i]] = (int32_t)(U - 1);
  }
  seg[U] = (int32_t)n;
  for (int64_t u = 0; u < U; ++u) counts[u] = seg[u + 1] - seg[u];
  free(ka); free(kb); free(pa); free(pb);
  return U;
}

/* ------------------------------------------------------------------ SPEC 'Shard function', 'A2 partition'
 * bases[T + 1] = exclusive prefix sum of rows_t; the shard of a fused key is (key - bases[t]) mod S. */
static int table_of(const int64_t* bases, int T, int64_t key) {
  int lo = 0, hi = T; /* last t with bases[t] <= key */
  while (hi - lo > 1) {
    const int mid = (lo + hi) / 2;
    if (bases[mid] <= key) lo = mid; else hi = mid;
  }
  return lo;
}

void edlo_partition_mod(const int64_t* keys, int64_t n, const int64_t* bases, int T, int S, int64_t* keys_out,
                        int32_t* offsets, int32_t* perm) {
  int32_t* shard = (int32_t*)malloc((size_t)(n > 0 ? n : 1) * 4);
  int64_t* fill = (int64_t*)calloc((size_t)S + 1, 8);
  for (int64_t i = 0; i < n; ++i) {
    const int t = table_of(bases, T, keys[i]);
    shard[i] = (int32_t)((keys[i] - bases[t]) % S);
    fill[shard[i] + 1]++;
  }
  for (int s = 0; s < S; ++s) fill[s + 1] += fill[s];
  for (int s = 0; s <= S; ++s) offsets[s] = (int32_t)fill[s];
  for (int64_t i = 0; i < n; ++i) { /* stable: ascending input position inside each group */
    const int64_t dst = fill[shard[i]]++;
    keys_out[dst] = keys[i];
    perm[dst] = (int32_t)i;
  }
  free(shard); free(fill);
}

/* ------------------------------------------------------------------ SPEC 'A4 gradient reduce'
 * rows of a segment added in stable sort order, in float64, rounded once to fp32 */
void edlo_segment_sum_f64(const float* grad, int dim, const int32_t* order, const int32_t* seg, int64_t U, float* out) {
  double* acc = (double*)malloc((size_t)dim * 8);
  for (int64_t u = 0; u < U; ++u) {
    for (int c = 0; c < dim; ++c) acc[c] = 0.0;
    for (int32_t i = seg[u]; i < seg[u + 1]; ++i) {
      const float* g = grad + (int64_t)order[i] * dim;
      if (i + 8 < seg[U]) prefetch_row(grad + (int64_t)order[i + 8] * dim, dim);
      acc_add(acc, g, dim);
    }
    acc_round(acc, out + u * dim, dim);
  }
  free(acc);
}

/* ------------------------------------------------------------------ SPEC 'A5 optimizer', 'A5 operation order' */
EDLO_CLONES void edlo_adagrad_rows(float* restrict p, float* restrict acc, const float* restrict g, int64_t n, float lr, float eps) {
  for (int64_t i = 0; i < n; ++i) {
    const float gg = g[i] * g[i];
    const float a = acc[i] + gg;
    const float num = lr * g[i];
    const float den = sqrtf(a) + eps;
    acc[i] = a;
    p[i] = p[i] - num / den;
  }
}

EDLO_CLONES void edlo_adam_rows(float* restrict p, float* restrict m, float* restrict v, const float* restrict g, int64_t n, float lr_t, float beta1, float beta2,
                    float eps) {
  const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
  for (int64_t i = 0; i < n; ++i) {
    const float m1 = beta1 * m[i];
    const float m2 = omb1 * g[i];
    const float mn = m1 + m2;
    const float v1 = beta2 * v[i];
    const float v2 = (omb2 * g[i]) * g[i];
    const float vn = v1 + v2;
    const float num = lr_t * mn;
    const float den = sqrtf(vn) + eps;
    m[i] = mn;
    v[i] = vn;
    p[i] = p[i] - num / den;
  }
}

float edlo_adam_lr_t(double lr, double beta1, double beta2, int64_t step) {
  return (float)(lr * sqrt(1.0 - pow(beta2, (double)step)) / (1.0 - pow(beta1, (double)step)));
}

/* ------------------------------------------------------------------ one PS shard (S == 1), lazily materialised
 * Per table: an open-addressing hash map id -> row number and growing row / slot stores.  A row exists
 * once it has received an update (SPEC 'Materialised row'); a lookup of any other row computes the pure
 * initialiser and stores nothing. */
typedef struct {
  int64_t* hkeys;  /* -1 = empty */
  int32_t* hrows;
  int64_t hcap;    /* power of two */
  int64_t n, cap;  /* rows stored / allocated */
  int64_t* ids;    /* id of row r */
  float *w, *s0, *s1;
} edlo_table;

typedef struct {
  int T, dim, kind;
  double lr, beta1, beta2;
  float eps, init_accum;
  uint64_t seed;
  int64_t step;
  int64_t* rows;
  edlo_table* tab;
} edlo_shard;

static uint64_t hash_id(int64_t id) { return edlo_splitmix64((uint64_t)id * 0x9E3779B97F4A7C15ull); }

static int64_t tab_find(const edlo_table* tb, int64_t id) {
  if (!tb->hcap) return -1;
  uint64_t h = hash_id(id) & (uint64_t)(tb->hcap - 1);
  while (tb->hkeys[h] != -1) {
    if (tb->hkeys[h] == id) return tb->hrows[h];
    h = (h + 1) & (uint64_t)(tb->hcap - 1);
  }
  return -1;
}

static void tab_rehash(edlo_table* tb, int64_t hcap) {
  int64_t* ok = tb->hkeys; int32_t* orow = tb->hrows; const int64_t ocap = tb->hcap;
  tb->hkeys = (int64_t*)malloc((size_t)hcap * 8);
  tb->hrows = (int32_t*)malloc((size_t)hcap * 4);
  tb->hcap = hcap;
  for (int64_t i = 0; i < hcap; ++i) tb->hkeys[i] = -1;
  for (int64_t i = 0; i < ocap; ++i)
    if (ok[i] != -1) {
      uint64_t h = hash_id(ok[i]) & (uint64_t)(hcap - 1);
      while (tb->hkeys[h] != -1) h = (h + 1) & (uint64_t)(hcap - 1);
      tb->hkeys[h] = ok[i]; tb->hrows[h] = orow[i];
    }
  free(ok); free(orow);
}

static int64_t tab_insert(edlo_shard* sh, int t, int64_t id) {
  edlo_table* tb = &sh->tab[t];
  const int dim = sh->dim;
  if ((tb->n + 1) * 2 > tb->hcap) tab_rehash(tb, tb->hcap ? tb->hcap * 2 : 1024);
  if (tb->n == tb->cap) {
    tb->cap = tb->cap ? tb->cap * 2 : 512;
    tb->ids = (int64_t*)realloc(tb->ids, (size_t)tb->cap * 8);
    tb->w = (float*)realloc(tb->w, (size_t)tb->cap * dim * 4);
    tb->s0 = (float*)realloc(tb->s0, (size_t)tb->cap * dim * 4);
    if (sh->kind == EDLO_ADAM) tb->s1 = (float*)realloc(tb->s1, (size_t)tb->cap * dim * 4);
  }
  const int64_t r = tb->n++;
  tb->ids[r] = id;
  init_row(sh->seed, t, id, dim, tb->w + r * dim);
  const float s0 = sh->kind == EDLO_ADAGRAD ? sh->init_accum : 0.0f;
  for (int c = 0; c < dim; ++c) tb->s0[r * dim + c] = s0;
  if (sh->kind == EDLO_ADAM) for (int c = 0; c < dim; ++c) tb->s1[r * dim + c] = 0.0f;
  uint64_t h = hash_id(id) & (uint64_t)(tb->hcap - 1);
  while (tb->hkeys[h] != -1) h = (h + 1) & (uint64_t)(tb->hcap - 1);
  tb->hkeys[h] = id; tb->hrows[h] = (int32_t)r;
  return r;
}

void* edlo_shard_create(const int64_t* rows, int T, int dim, int kind, double lr, double eps, double init_accum,
                        double beta1, double beta2, uint64_t seed) {
  edlo_shard* sh = (edlo_shard*)calloc(1, sizeof(edlo_shard));
  sh->T = T; sh->dim = dim; sh->kind = kind; sh->lr = lr; sh->beta1 = beta1; sh->beta2 = beta2;
  sh->eps = (float)eps; sh->init_accum = (float)init_accum; sh->seed = seed;
  sh->rows = (int64_t*)malloc((size_t)T * 8);
  memcpy(sh->rows, rows, (size_t)T * 8);
  sh->tab = (edlo_table*)calloc((size_t)T, sizeof(edlo_table));
  return sh;
}

void edlo_shard_destroy(void* p) {
  edlo_shard* sh = (edlo_shard*)p;
  if (!sh) return;
  for (int t = 0; t < sh->T; ++t) {
    edlo_table* tb = &sh->tab[t];
    free(tb->hkeys); free(tb->hrows); free(tb->ids); free(tb->w); free(tb->s0); free(tb->s1);
  }
  free(sh->tab); free(sh->rows); free(sh);
}

int64_t edlo_shard_step_count(const void* p) { return ((const edlo_shard*)p)->step; }

int64_t edlo_shard_rows(const void* p) {
  const edlo_shard* sh = (const edlo_shard*)p;
  int64_t n = 0;
  for (int t = 0; t < sh->T; ++t) n += sh->tab[t].n;
  return n;
}

/* One table of one step: A1 on the table's ids, A3 (rows as they are BEFORE this step's update) + expand
 * into emb[b, t, :], A4, A5 with insert-if-absent.  Returns the number of unique ids. */
static int64_t step_table(edlo_shard* sh, int t, const int64_t* ids, int64_t B, int F, const float* grad, float* emb,
                          float rate) {
  const int dim = sh->dim;
  int64_t* col = (int64_t*)malloc((size_t)B * 8);
  int64_t* uniq = (int64_t*)malloc((size_t)B * 8);
  int32_t* inverse = (int32_t*)malloc((size_t)B * 4);
  int32_t* counts = (int32_t*)malloc((size_t)B * 4);
  int32_t* order = (int32_t*)malloc((size_t)B * 4);
  int32_t* seg = (int32_t*)malloc((size_t)(B + 1) * 4);
  for (int64_t b = 0; b < B; ++b) col[b] = ids[b * F + t];
  const int64_t U = edlo_dedup(col, B, uniq, inverse, counts, order, seg);
  float* rows_u = (float*)malloc((size_t)(U > 0 ? U : 1) * dim * 4);
  edlo_table* tb = &sh->tab[t];
  for (int64_t u = 0; u < U; ++u) {
    const int64_t r = tab_find(tb, uniq[u]);
    if (r >= 0) memcpy(rows_u + u * dim, tb->w + r * dim, (size_t)dim * 4);
    else init_row(sh->seed, t, uniq[u], dim, rows_u + u * dim);
  }
  if (emb)
    for (int64_t b = 0; b < B; ++b)
      memcpy(emb + (b * F + t) * dim, rows_u + (int64_t)inverse[b] * dim, (size_t)dim * 4);
  /* A4 over the table's column of the [B, F, dim] gradient, then A5 row by row */
  double* acc = (double*)malloc((size_t)dim * 8);
  float* g = (float*)malloc((size_t)dim * 4);
  for (int64_t u = 0; u < U; ++u) {
    for (int c = 0; c < dim; ++c) acc[c] = 0.0;
    for (int32_t i = seg[u]; i < seg[u + 1]; ++i) {
      const float* gr = grad + ((int64_t)order[i] * F + t) * dim;
      if (i + 8 < seg[U]) prefetch_row(grad + ((int64_t)order[i + 8] * F + t) * dim, dim);
      acc_add(acc, gr, dim);
    }
    acc_round(acc, g, dim);
    int64_t r = tab_find(tb, uniq[u]);
    if (r < 0) r = tab_insert(sh, t, uniq[u]);
    if (sh->kind == EDLO_ADAGRAD)
      edlo_adagrad_rows(tb->w + r * dim, tb->s0 + r * dim, g, dim, rate, sh->eps);
    else
      edlo_adam_rows(tb->w + r * dim, tb->s0 + r * dim, tb->s1 + r * dim, g, dim, rate, (float)sh->beta1,
                     (float)sh->beta2, sh->eps);
  }
  free(acc); free(g); free(rows_u); free(col); free(uniq); free(inverse); free(counts); free(order); free(seg);
  return U;
}

/* ids[B, F] per-table ids (field f reads table f), grad[B * F, dim], emb[B * F, dim] (nullable).
 * Tables are independent, so n_threads threads take them one at a time from a shared counter.
 * -> sum of unique ids per table. */
typedef struct {
  edlo_shard* sh;
  const int64_t* ids;
  int64_t B;
  int F;
  const float* grad;
  float* emb;
  float rate;
  int next;        /* next table to hand out */
  int64_t total;
} step_job;

static void* step_worker(void* arg) {
  step_job* job = (step_job*)arg;
  for (;;) {
    const int t = __sync_fetch_and_add(&job->next, 1);
    if (t >= job->F) return NULL;
    const int64_t u = step_table(job->sh, t, job->ids, job->B, job->F, job->grad, job->emb, job->rate);
    __sync_fetch_and_add(&job->total, u);
  }
}

int64_t edlo_shard_step(void* p, const int64_t* ids, int64_t B, int F, const float* grad, float* emb, int n_threads) {
  edlo_shard* sh = (edlo_shard*)p;
  sh->step += 1;
  step_job job;
  job.sh = sh; job.ids = ids; job.B = B; job.F = F; job.grad = grad; job.emb = emb; job.next = 0; job.total = 0;
  job.rate = sh->kind == EDLO_ADAGRAD ? (float)sh->lr : edlo_adam_lr_t(sh->lr, sh->beta1, sh->beta2, sh->step);
  if (n_threads > F) n_threads = F;
  if (n_threads <= 1) {
    step_worker(&job);
    return job.total;
  }
  pthread_t* th = (pthread_t*)malloc((size_t)n_threads * sizeof(pthread_t));
  int started = 0;
  for (int i = 0; i < n_threads - 1; ++i)
    if (pthread_create(&th[started], NULL, step_worker, &job) == 0) ++started;
  step_worker(&job);              /* the calling thread works too (and alone, if no thread could start) */
  for (int i = 0; i < started; ++i) pthread_join(th[i], NULL);
  free(th);
  return job.total;
}

/* SPEC 'A6': materialised rows ascending by fused key (table-major, ids ascending inside a table).
 * keys_out[n], w_out / s0_out / s1_out [n, dim] with n = edlo_shard_rows(); s1_out nullable. */
void edlo_shard_export(const void* p, int64_t* keys_out, float* w_out, float* s0_out, float* s1_out) {
  const edlo_shard* sh = (const edlo_shard*)p;
  const int dim = sh->dim;
  int64_t base = 0, o = 0;
  for (int t = 0; t < sh->T; ++t) {
    const edlo_table* tb = &sh->tab[t];
    const int64_t n = tb->n;
    if (n) {
      int64_t* uniq = (int64_t*)malloc((size_t)n * 8);
      int32_t* inverse = (int32_t*)malloc((size_t)n * 4);
      int32_t* counts = (int32_t*)malloc((size_t)n * 4);
      int32_t* order = (int32_t*)malloc((size_t)n * 4);
      int32_t* seg = (int32_t*)malloc((size_t)(n + 1) * 4);
      edlo_dedup(tb->ids, n, uniq, inverse, counts, order, seg);
      for (int64_t i = 0; i < n; ++i, ++o) {
        const int64_t r = order[i];
        keys_out[o] = base + tb->ids[r];
        memcpy(w_out + o * dim, tb->w + r * dim, (size_t)dim * 4);
        memcpy(s0_out + o * dim, tb->s0 + r * dim, (size_t)dim * 4);
        if (s1_out && tb->s1) memcpy(s1_out + o * dim, tb->s1 + r * dim, (size_t)dim * 4);
      }
      free(uniq); free(inverse); free(counts); free(order); free(seg);
    }
    base += sh->rows[t];
  }
}

/* ------------------------------------------------------------------ SPEC 'A6 re-shard', 'Checkpoint file'
 * Writes the file shard `new_shard` of `new_S` would hold after a re-shard of this (single) shard: the
 * materialised records with id mod new_S == new_shard, per table, ids ascending.  Little-endian:
 * "EDLS", version u32, S u32, shard u32, T u32, step u64, optimizer u32; per table t u32, D u32, n u64,
 * ids[n] i64, rows[n * D] f32, slot0[n * D] f32 (, slot1[n * D] f32 for Adam).  -> 0, or -1 on an I/O error. */
static int put(FILE* f, const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }

int edlo_shard_write_checkpoint(const void* p, const char* path, int new_S, int new_shard) {
  const edlo_shard* sh = (const edlo_shard*)p;
  const int dim = sh->dim;
  FILE* f = fopen(path, "wb");
  if (!f) return -1;
  const uint32_t head[4] = {1u, (uint32_t)new_S, (uint32_t)new_shard, (uint32_t)sh->T};
  const uint64_t step = (uint64_t)sh->step;
  const uint32_t kind = (uint32_t)sh->kind;
  int ok = put(f, "EDLS", 4) && put(f, head, sizeof head) && put(f, &step, 8) && put(f, &kind, 4);
  for (int t = 0; t < sh->T && ok; ++t) {
    const edlo_table* tb = &sh->tab[t];
    const int64_t n = tb->n;
    int64_t* uniq = (int64_t*)malloc((size_t)(n + 1) * 8);
    int32_t* inverse = (int32_t*)malloc((size_t)(n + 1) * 4);
    int32_t* counts = (int32_t*)malloc((size_t)(n + 1) * 4);
    int32_t* order = (int32_t*)malloc((size_t)(n + 1) * 4);
    int32_t* seg = (int32_t*)malloc((size_t)(n + 2) * 4);
    edlo_dedup(tb->ids, n, uniq, inverse, counts, order, seg); /* order = rows by ascending id */
    uint64_t m = 0;
    for (int64_t i = 0; i < n; ++i) m += (tb->ids[order[i]] % new_S) == new_shard;
    const uint32_t th[2] = {(uint32_t)t, (uint32_t)dim};
    ok = ok && put(f, th, sizeof th) && put(f, &m, 8);
    for (int64_t i = 0; i < n && ok; ++i)
      if ((tb->ids[order[i]] % new_S) == new_shard) ok = put(f, &tb->ids[order[i]], 8);
    const float* arrays[3] = {tb->w, tb->s0, sh->kind == EDLO_ADAM ? tb->s1 : NULL};
    for (int a = 0; a < 3 && ok; ++a) {
      if (!arrays[a]) continue;
      for (int64_t i = 0; i < n && ok; ++i)
        if ((tb->ids[order[i]] % new_S) == new_shard) ok = put(f, arrays[a] + (int64_t)order[i] * dim, (size_t)dim * 4);
    }
    free(uniq); free(inverse); free(counts); free(order); free(seg);
  }
  return (fclose(f) == 0 && ok) ? 0 : -1;
}
