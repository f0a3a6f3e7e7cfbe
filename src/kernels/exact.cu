// Direct fp64 brute force on the GPU: every (query, datapoint) distance with the
// reference's rounding, then an exact sort.  This is the reference's algorithm
// as written (engine.cpp:233-257) moved onto the device; the engine uses it for
// problems too small to be worth a tensor-core pass, for k beyond the fused
// epilogue's list capacity, and for queries the filter could not certify.
//
// Distances are written in descending-id order and sorted with a stable radix
// sort, which yields exactly the (distance asc, id desc) order of the spec.
#include <cub/cub.cuh>

#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

namespace {

constexpr int EX_XR = 64;    // datapoints per block tile
constexpr int EX_QB = 16;    // queries per block tile
constexpr int EX_CH = 32;    // columns per step
constexpr int EX_THREADS = 256;
constexpr int EX_SEG_LIMIT = 16384;  // segments up to this length use the segmented sort

__global__ void __launch_bounds__(EX_THREADS)
exact_dist_kernel(const double *__restrict__ q_attrs, const double *__restrict__ x_attrs,
                  const int *__restrict__ query_rows, int qbase, int count, int n, int a,
                  double *__restrict__ dist_out, int *__restrict__ id_out) {
  __shared__ double xs[EX_XR][EX_CH + 1];
  __shared__ double qs[EX_QB][EX_CH + 1];
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * EX_XR;
  const int b0 = blockIdx.y * EX_QB;
  const int xr = tid % EX_XR;
  const int qg = tid / EX_XR;  // 4 groups of 4 queries
  double acc[4] = {0.0, 0.0, 0.0, 0.0};

  for (int c0 = 0; c0 < a; c0 += EX_CH) {
    for (int i = tid; i < EX_XR * EX_CH; i += EX_THREADS) {
      const int r = i / EX_CH, c = i - r * EX_CH;
      const int row = x0 + r, col = c0 + c;
      xs[r][c] = (row < n && col < a) ? x_attrs[static_cast<long long>(row) * a + col] : 0.0;
    }
    for (int i = tid; i < EX_QB * EX_CH; i += EX_THREADS) {
      const int r = i / EX_CH, c = i - r * EX_CH;
      const int b = b0 + r, col = c0 + c;
      double v = 0.0;
      if (b < count && col < a) {
        const int qrow = query_rows ? query_rows[qbase + b] : qbase + b;
        v = q_attrs[static_cast<long long>(qrow) * a + col];
      }
      qs[r][c] = v;
    }
    __syncthreads();
#pragma unroll 4
    for (int i = 0; i < EX_CH; ++i) {
      const double xv = xs[xr][i];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const double d = __dsub_rn(qs[qg * 4 + j][i], xv);
        acc[j] = __dadd_rn(acc[j], __dmul_rn(d, d));  // zero padding adds +0.0 exactly
      }
    }
    __syncthreads();
  }
  const int row = x0 + xr;
  if (row < n) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int b = b0 + qg * 4 + j;
      if (b < count) {
        const long long pos = static_cast<long long>(b) * n + (n - 1 - row);  // descending id
        dist_out[pos] = acc[j];
        id_out[pos] = row;
      }
    }
  }
}

__global__ void exact_emit_kernel(const double *__restrict__ sorted_dist, const int *__restrict__ sorted_id,
                                  const int *__restrict__ x_labels, const int *__restrict__ kq,
                                  const int *__restrict__ query_rows, int qbase, int count, int n, int id_base,
                                  int kmax, double *out_dist, int *out_id, int *out_label) {
  const int b = blockIdx.x;
  if (b >= count) return;
  const int qrow = query_rows ? query_rows[qbase + b] : qbase + b;
  int k = kq[qrow];
  k = k < 0 ? 0 : (k > kmax ? kmax : k);
  for (int i = threadIdx.x; i < kmax; i += blockDim.x) {
    const long long o = static_cast<long long>(qrow) * kmax + i;
    if (i < k && i < n) {
      const int id = sorted_id[static_cast<long long>(b) * n + i];
      out_dist[o] = sorted_dist[static_cast<long long>(b) * n + i];
      out_id[o] = id_base + id;
      out_label[o] = x_labels[id];
    } else {
      out_dist[o] = 1e18;
      out_id[o] = -1;
      out_label[o] = -1;
    }
  }
}

__global__ void fill_offsets_kernel(int *off, int count, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= count) off[i] = i * n;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

size_t sort_temp_bytes(int n, int batch) {
  size_t seg = 0, one = 0;
  cub::DeviceSegmentedRadixSort::SortPairs(nullptr, seg, (const double *)nullptr, (double *)nullptr,
                                           (const int *)nullptr, (int *)nullptr, batch * n, batch,
                                           (const int *)nullptr, (const int *)nullptr);
  cub::DeviceRadixSort::SortPairs(nullptr, one, (const double *)nullptr, (double *)nullptr,
                                  (const int *)nullptr, (int *)nullptr, n);
  return seg > one ? seg : one;
}

}  // namespace

size_t exact_workspace_bytes(int n, int batch) {
  if (n <= 0 || batch <= 0) return 256;
  const size_t items = static_cast<size_t>(n) * batch;
  return align_up(items * 8, 256) * 2 + align_up(items * 4, 256) * 2 + align_up((batch + 1) * 4, 256) +
         align_up(sort_temp_bytes(n, batch), 256) + 256;
}

int launch_exact(const ExactArgs &args, cudaStream_t s) {
  if (args.count <= 0) return 0;
  DMLP_REQUIRE(args.kmax > 0, "exact: kmax must be positive");
  if (args.n <= 0) {
    exact_emit_kernel<<<args.count, 64, 0, s>>>(nullptr, nullptr, args.x_labels, args.k, args.query_rows, 0,
                                                args.count, 0, args.id_base, args.kmax, args.out_dist, args.out_id,
                                                args.out_label);
    DMLP_CUDA_OK(cudaGetLastError());
    return 0;
  }
  DMLP_REQUIRE(static_cast<long long>(args.n) * EX_QB < (1LL << 31), "exact: shard too large for 32-bit offsets");
  // largest batch the workspace supports
  int batch = args.count;
  while (batch > 1 && exact_workspace_bytes(args.n, batch) > args.workspace_bytes) batch = (batch + 1) / 2;
  while (static_cast<long long>(batch) * args.n >= (1LL << 31)) batch = (batch + 1) / 2;
  DMLP_REQUIRE(exact_workspace_bytes(args.n, batch) <= args.workspace_bytes,
               "exact: workspace too small for even one query");
  const size_t items = static_cast<size_t>(args.n) * batch;
  unsigned char *w = static_cast<unsigned char *>(args.workspace);
  double *d_in = reinterpret_cast<double *>(w); w += align_up(items * 8, 256);
  double *d_out = reinterpret_cast<double *>(w); w += align_up(items * 8, 256);
  int *i_in = reinterpret_cast<int *>(w); w += align_up(items * 4, 256);
  int *i_out = reinterpret_cast<int *>(w); w += align_up(items * 4, 256);
  int *offs = reinterpret_cast<int *>(w); w += align_up((batch + 1) * 4, 256);
  void *temp = w;
  size_t temp_bytes = sort_temp_bytes(args.n, batch);

  for (int base = 0; base < args.count; base += batch) {
    const int cnt = args.count - base < batch ? args.count - base : batch;
    dim3 grid((args.n + EX_XR - 1) / EX_XR, (cnt + EX_QB - 1) / EX_QB);
    exact_dist_kernel<<<grid, EX_THREADS, 0, s>>>(args.q_attrs, args.x_attrs, args.query_rows, base, cnt, args.n,
                                                  args.a, d_in, i_in);
    DMLP_CUDA_OK(cudaGetLastError());
    if (args.n <= EX_SEG_LIMIT) {
      fill_offsets_kernel<<<(cnt + 256) / 256, 256, 0, s>>>(offs, cnt, args.n);
      size_t tb = temp_bytes;
      DMLP_CUDA_OK(cub::DeviceSegmentedRadixSort::SortPairs(temp, tb, d_in, d_out, i_in, i_out, cnt * args.n, cnt,
                                                            offs, offs + 1, 0, 64, s));
    } else {
      for (int b = 0; b < cnt; ++b) {
        size_t tb = temp_bytes;
        const size_t o = static_cast<size_t>(b) * args.n;
        DMLP_CUDA_OK(cub::DeviceRadixSort::SortPairs(temp, tb, d_in + o, d_out + o, i_in + o, i_out + o, args.n, 0,
                                                     64, s));
      }
    }
    exact_emit_kernel<<<cnt, 128, 0, s>>>(d_out, i_out, args.x_labels, args.k, args.query_rows, base, cnt, args.n,
                                          args.id_base, args.kmax, args.out_dist, args.out_id, args.out_label);
    DMLP_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

}  // namespace k
}  // namespace dmlp
