// K5 (owner side) + K6: merge the per-shard exact k-lists of each query, vote,
// order and fold the checksum.
//
// Reference counterpart: row 0 of each grid column re-assembles the gathered
// lists and selects k again (engine.cpp:289-309), then votes, sorts and reports
// (engine.cpp:320-347, common.cpp:59-68).  Each source list arrives sorted by
// (distance asc, id desc) and padded with (1e18, -1) sentinels, so selection is
// a rank computation over at most sources * kmax candidates -- the device form
// of the staff baselines' "merge two sorted k-lists" reduction operator.
#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

namespace {

constexpr int FIN_WARPS = 4;
constexpr int FIN_MAXC = 1024;  // candidates per query handled on the device

__global__ void __launch_bounds__(FIN_WARPS * 32) finalize_kernel(const FinalizeArgs a) {
  extern __shared__ __align__(16) unsigned char fin_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * FIN_WARPS + warp;
  if (q >= a.q) return;
  // per-warp staging: candidates then the final list
  double *c_dist = reinterpret_cast<double *>(fin_smem) + static_cast<size_t>(warp) * FIN_MAXC;
  int *c_id = reinterpret_cast<int *>(fin_smem + sizeof(double) * FIN_WARPS * FIN_MAXC) + warp * 2 * FIN_MAXC;
  int *c_label = c_id + FIN_MAXC;
  double *f_dist = reinterpret_cast<double *>(fin_smem + (sizeof(double) + 2 * sizeof(int)) * FIN_WARPS * FIN_MAXC) +
                   static_cast<size_t>(warp) * FIN_MAXC;
  int *f_id = reinterpret_cast<int *>(fin_smem + (2 * sizeof(double) + 2 * sizeof(int)) * FIN_WARPS * FIN_MAXC) +
              warp * 2 * FIN_MAXC;
  int *f_label = f_id + FIN_MAXC;

  int kq = a.k[q];
  kq = kq < 0 ? 0 : (kq > a.kmax ? a.kmax : kq);
  const int total = a.sources * kq;
  for (int i = lane; i < total; i += 32) {
    const int src = i / kq, j = i - src * kq;
    const long long off = src * a.source_stride + static_cast<long long>(q) * a.kmax + j;
    c_dist[i] = a.dist[off];
    c_id[i] = a.id[off];
    c_label[i] = a.label[off];
  }
  for (int i = lane; i < kq; i += 32) { f_dist[i] = 1e18; f_id[i] = -1; f_label[i] = -1; }
  __syncwarp();
  for (int i = lane; i < total; i += 32) {
    const double d = c_dist[i];
    const int id = c_id[i];
    if (id < 0) continue;  // sentinels only fill what real neighbours leave empty
    int rank = 0;
    for (int j = 0; j < total; ++j) {
      const double dj = c_dist[j];
      rank += (dj < d || (dj == d && c_id[j] > id)) ? 1 : 0;
    }
    if (rank < kq) { f_dist[rank] = d; f_id[rank] = id; f_label[rank] = c_label[i]; }
  }
  __syncwarp();

  // majority vote, ties to the larger label, -1 when nothing voted
  unsigned long long best = 0;
  for (int i = lane; i < kq; i += 32) {
    if (f_id[i] < 0) continue;
    const int lab = f_label[i];
    int count = 0;
    for (int j = 0; j < kq; ++j) count += (f_id[j] >= 0 && f_label[j] == lab) ? 1 : 0;
    const unsigned long long key =
        (static_cast<unsigned long long>(count) << 32) | (static_cast<uint32_t>(lab) ^ 0x80000000u);
    best = key > best ? key : best;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    best = other > best ? other : best;
  }
  const int label = best == 0 ? -1 : static_cast<int>(static_cast<uint32_t>(best & 0xFFFFFFFFu) ^ 0x80000000u);

  if (lane == 0) {
    unsigned long long h = 1469598103934665603ULL;
    h ^= static_cast<unsigned long long>(static_cast<long long>(label));
    h *= 1099511628211ULL;
    for (int i = 0; i < kq; ++i) {
      h ^= static_cast<unsigned long long>(static_cast<long long>(f_id[i] + 1));
      h *= 1099511628211ULL;
    }
    a.out_label[q] = label;
    a.out_checksum[q] = h;
  }
  if (a.out_id != nullptr)
    for (int i = lane; i < a.kmax; i += 32) a.out_id[static_cast<long long>(q) * a.kmax + i] = i < kq ? f_id[i] : -1;
  if (a.out_dist != nullptr)
    for (int i = lane; i < a.kmax; i += 32)
      a.out_dist[static_cast<long long>(q) * a.kmax + i] = i < kq ? f_dist[i] : 1e18;
}

}  // namespace

int launch_finalize(const FinalizeArgs &args, cudaStream_t s) {
  if (args.q <= 0) return 0;
  DMLP_REQUIRE(args.sources >= 1 && args.kmax >= 1, "finalize: need at least one source and kmax >= 1");
  DMLP_REQUIRE(static_cast<long long>(args.sources) * args.kmax <= FIN_MAXC,
               "finalize: sources * kmax exceeds the device merge capacity (1024)");
  const size_t smem = (2 * sizeof(double) + 4 * sizeof(int)) * FIN_WARPS * FIN_MAXC;
  DMLP_CUDA_OK(cudaFuncSetAttribute(finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  const int blocks = (args.q + FIN_WARPS - 1) / FIN_WARPS;
  finalize_kernel<<<blocks, FIN_WARPS * 32, smem, s>>>(args);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
