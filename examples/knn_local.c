/* A whole k-NN job on one GPU through the C ABI (include/dmlp_knn.h), in plain C.
 *
 * The library never allocates: this program owns every buffer.  It generates a problem on the
 * device with the reference generator's distribution (dmlp_synth_*), runs the local pipeline
 * (pack -> tcgen05 filter GEMM with the top-k fused in -> certified exact fp64 re-rank), repairs
 * whatever the filter could not certify with the direct kernel, votes and checksums, and prints
 * the reference's release-mode lines ("Query i checksum: c").
 *
 *   make examples && ./examples/knn_local [num_data num_queries num_attrs kmax]
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "dmlp_knn.h"

#define CUDA_OK(call)                                                                     \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));          \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)
#define DMLP_OK(call)                                                                     \
  do {                                                                                    \
    if ((call) != 0) {                                                                    \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, dmlp_last_error());              \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

int main(int argc, char **argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 100000, q = argc > 2 ? atoi(argv[2]) : 2000;
  const int a = argc > 3 ? atoi(argv[3]) : 64, kmax = argc > 4 ? atoi(argv[4]) : 16;
  int sms = 0, cc_major = 0, cc_minor = 0;
  DMLP_OK(dmlp_device_info(&sms, &cc_major, &cc_minor));
  if (dmlp_fused_supported(kmax, n) < 1) {
    fprintf(stderr, "k = %d of n = %d points: use dmlp_exact_ex (the direct kernel) for this one\n", kmax, n);
    return 1;
  }

  cudaStream_t st;
  CUDA_OK(cudaStreamCreate(&st));
  double *x, *qa, *dist, *out_dist;
  int *labels, *k, *id, *lab, *flags, *nflag, *out_label, *out_id;
  uint64_t *out_sum;
  void *ws;
  const size_t ws_bytes = dmlp_knn_local_workspace_bytes(q, n, a, sms, kmax);
  const size_t slots = (size_t)q * (size_t)kmax;
  CUDA_OK(cudaMalloc((void **)&x, sizeof(double) * (size_t)n * a));
  CUDA_OK(cudaMalloc((void **)&labels, sizeof(int) * (size_t)n));
  CUDA_OK(cudaMalloc((void **)&qa, sizeof(double) * (size_t)q * a));
  CUDA_OK(cudaMalloc((void **)&k, sizeof(int) * (size_t)q));
  CUDA_OK(cudaMalloc((void **)&dist, sizeof(double) * slots));
  CUDA_OK(cudaMalloc((void **)&id, sizeof(int) * slots));
  CUDA_OK(cudaMalloc((void **)&lab, sizeof(int) * slots));
  CUDA_OK(cudaMalloc((void **)&flags, sizeof(int) * (size_t)q));
  CUDA_OK(cudaMalloc((void **)&nflag, sizeof(int)));
  CUDA_OK(cudaMalloc((void **)&out_label, sizeof(int) * (size_t)q));
  CUDA_OK(cudaMalloc((void **)&out_sum, sizeof(uint64_t) * (size_t)q));
  CUDA_OK(cudaMalloc((void **)&out_id, sizeof(int) * slots));
  CUDA_OK(cudaMalloc((void **)&out_dist, sizeof(double) * slots));
  CUDA_OK(cudaMalloc(&ws, ws_bytes));

  /* the generator's problem: values in [-100, 100] with six decimals, 10 labels, k in [1, kmax] */
  DMLP_OK(dmlp_synth_attrs(x, n, a, 0, -100.0, 100.0, 42, 0, st));
  DMLP_OK(dmlp_synth_labels(labels, n, 0, 10, 42, st));
  DMLP_OK(dmlp_synth_attrs(qa, q, a, 0, -100.0, 100.0, 42, 2, st));
  DMLP_OK(dmlp_synth_k(k, q, 0, 1, kmax, 42, st));

  CUDA_OK(cudaMemsetAsync(nflag, 0, sizeof(int), st));
  DMLP_OK(dmlp_knn_local(qa, k, q, x, labels, n, a, /*id_base=*/0, kmax, /*range=*/NULL, ws, ws_bytes, dist, id, lab,
                         flags, nflag, st));
  int flagged = 0;
  CUDA_OK(cudaMemcpyAsync(&flagged, nflag, sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  if (flagged > 0) {
    /* queries the filter could not certify: recompute them directly (rare; none on this distribution).
     * The scratch of knn_local is free again, so the direct kernel can have it. */
    int *host_flags = (int *)malloc(sizeof(int) * (size_t)q), *rows_host = (int *)malloc(sizeof(int) * (size_t)flagged);
    int *rows;
    int cnt = 0;
    CUDA_OK(cudaMemcpy(host_flags, flags, sizeof(int) * (size_t)q, cudaMemcpyDeviceToHost));
    for (int i = 0; i < q && cnt < flagged; ++i)
      if (host_flags[i]) rows_host[cnt++] = i;
    CUDA_OK(cudaMalloc((void **)&rows, sizeof(int) * (size_t)cnt));
    CUDA_OK(cudaMemcpy(rows, rows_host, sizeof(int) * (size_t)cnt, cudaMemcpyHostToDevice));
    const size_t need = dmlp_exact_workspace_bytes(n, cnt, kmax);
    void *ews = ws;
    if (need > ws_bytes) CUDA_OK(cudaMalloc(&ews, need));
    dmlp_exact_args e = {0};
    e.q_attrs = qa; e.x_attrs = x; e.x_labels = labels; e.k = k; e.query_rows = rows; e.count = cnt;
    e.n = n; e.a = a; e.id_base = 0; e.kmax = kmax; e.out_dist = dist; e.out_id = id; e.out_label = lab;
    e.workspace = ews; e.workspace_bytes = need > ws_bytes ? need : ws_bytes;
    DMLP_OK(dmlp_exact_ex(&e, st));
    CUDA_OK(cudaStreamSynchronize(st));
    free(host_flags);
    free(rows_host);
  }
  /* one source of lists: vote (ties to the larger label) + the reference's checksum */
  DMLP_OK(dmlp_finalize(dist, id, lab, /*source_stride=*/0, /*sources=*/1, q, kmax, k, out_label, out_sum, out_id,
                        out_dist, NULL, 0, st));
  uint64_t *sums = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)q);
  CUDA_OK(cudaMemcpyAsync(sums, out_sum, sizeof(uint64_t) * (size_t)q, cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  for (int i = 0; i < q && i < 5; ++i) printf("Query %d checksum: %llu\n", i, (unsigned long long)sums[i]);
  fprintf(stderr, "%d queries x %d points x %d attributes on %d SMs (sm_%d%d), %d flagged\n", q, n, a, sms, cc_major,
          cc_minor, flagged);
  free(sums);
  return 0;
}
