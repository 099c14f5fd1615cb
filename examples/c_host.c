/* A plain C host of libedlsparse.so: one dedup + segmented reduce + Adagrad update on a tiny batch,
 * checked against the same arithmetic done on the host.  Shows the drop-in boundary as a C program
 * would use it (raw device pointers, caller-owned workspace, error string), with no Python or torch.
 *
 *   gcc -std=c99 -I include -I /usr/local/cuda/include examples/c_host.c \
 *       -L easydl_b200 -ledlsparse -L /usr/local/cuda/lib64 -lcudart -lm -o /tmp/c_host
 *   LD_LIBRARY_PATH=easydl_b200 /tmp/c_host            (needs a GPU)
 */
#include <cuda_runtime_api.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "edlsparse.h"

#define N 12
#define DIM 4
#define ROWS 8

#define CK(call)                                                               \
  do {                                                                         \
    int rc_ = (call);                                                          \
    if (rc_ != 0) {                                                            \
      fprintf(stderr, "%s -> %d: %s\n", #call, rc_, edl_last_error_string()); \
      return 1;                                                                \
    }                                                                          \
  } while (0)
#define CU(call)                                                        \
  do {                                                                  \
    cudaError_t e_ = (call);                                            \
    if (e_ != cudaSuccess) {                                            \
      fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_));       \
      return 1;                                                         \
    }                                                                   \
  } while (0)

int main(void) {
  const int64_t ids[N] = {3, 1, 3, 7, 0, 1, 3, 5, 7, 7, 2, 3};
  float grad[N * DIM], slab[ROWS * DIM], acc[ROWS * DIM];
  for (int i = 0; i < N * DIM; ++i) grad[i] = 0.01f * (float)((i * 7) % 13 - 6);
  for (int i = 0; i < ROWS * DIM; ++i) { slab[i] = 0.5f - 0.01f * (float)i; acc[i] = 0.1f; }

  int64_t *d_ids, *d_uniq;
  int32_t *d_order, *d_seg, *d_first, *d_nu;
  float *d_grad, *d_slab, *d_acc;
  void *ws_u, *ws_s;
  const size_t wsu = edl_unique_workspace_bytes(N, 3, 0, 0);
  const size_t wss = edl_segreduce_workspace_bytes(N, N, DIM);
  CU(cudaMalloc((void**)&d_ids, sizeof ids));
  CU(cudaMalloc((void**)&d_uniq, N * sizeof(int64_t)));
  CU(cudaMalloc((void**)&d_order, N * sizeof(int32_t)));
  CU(cudaMalloc((void**)&d_seg, (N + 1) * sizeof(int32_t)));
  CU(cudaMalloc((void**)&d_first, N * sizeof(int32_t)));
  CU(cudaMalloc((void**)&d_nu, sizeof(int32_t)));
  CU(cudaMalloc((void**)&d_grad, sizeof grad));
  CU(cudaMalloc((void**)&d_slab, sizeof slab));
  CU(cudaMalloc((void**)&d_acc, sizeof acc));
  CU(cudaMalloc(&ws_u, wsu));
  CU(cudaMalloc(&ws_s, wss));
  CU(cudaMemcpy(d_ids, ids, sizeof ids, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d_grad, grad, sizeof grad, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d_slab, slab, sizeof slab, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d_acc, acc, sizeof acc, cudaMemcpyHostToDevice));

  /* A1: keys = ids (one table, 3 key bits), no inverse / counts wanted */
  CK(edl_unique(d_ids, N, NULL, NULL, 0, 3, 0, d_uniq, NULL, NULL, d_order, d_seg, d_first, d_nu, ws_u, wsu, NULL));
  /* A4 + A5: layout NULL = one shard, slab row == key */
  edl_opt_params opt;
  memset(&opt, 0, sizeof opt);
  opt.kind = EDL_OPT_ADAGRAD; opt.lr = 0.05f; opt.eps = 1e-7f;
  CK(edl_segreduce_update(d_slab, d_acc, NULL, DIM, 0 /* dense rows */, d_uniq, d_order, d_seg, d_first, d_nu, N, N, d_grad, 0, NULL, NULL,
                          &opt, NULL, ws_s, wss, NULL));
  CU(cudaDeviceSynchronize());

  float got[ROWS * DIM];
  int32_t nu = 0;
  CU(cudaMemcpy(got, d_slab, sizeof got, cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(&nu, d_nu, sizeof nu, cudaMemcpyDeviceToHost));

  /* the same on the host: float64 segment sum rounded once, then Adagrad op by op in fp32 */
  int bad = 0, seen = 0;
  for (int r = 0; r < ROWS; ++r) {
    int cnt = 0;
    double s[DIM] = {0, 0, 0, 0};
    for (int j = 0; j < N; ++j)
      if (ids[j] == r) { ++cnt; for (int c = 0; c < DIM; ++c) s[c] += (double)grad[j * DIM + c]; }
    seen += cnt > 0;
    for (int c = 0; c < DIM; ++c) {
      float want = slab[r * DIM + c];
      if (cnt) {
        volatile float g = (float)s[c];
        volatile float a = acc[r * DIM + c] + g * g;      /* volatile: keep every op a separate fp32 rounding */
        volatile float den = sqrtf(a) + opt.eps;
        volatile float step = opt.lr * g;
        volatile float upd = step / den;
        want = want - upd;
      }
      if (fabsf(got[r * DIM + c] - want) > 1e-6f * fabsf(want) + 1e-9f) ++bad;
    }
  }
  printf("unique ids: %d (expected %d); mismatching elements: %d\n", (int)nu, seen, bad);
  return (bad == 0 && nu == seen) ? 0 : 2;
}
