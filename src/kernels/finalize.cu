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
#include "merge.cuh"

namespace dmlp {
namespace k {

namespace {

constexpr int FIN_WARPS = 4;
constexpr int FIN_SMEM_K = 2048;  // merged lists up to this k are staged in shared memory, longer ones in the workspace

// One warp per query.  With one source the list is final as it stands; with several the
// entries are ranked against the other sources' lists by binary search (merge.cuh): no
// candidate staging and no limit on sources * kmax.
__global__ void __launch_bounds__(FIN_WARPS * 32) finalize_kernel(const FinalizeArgs a) {
  extern __shared__ __align__(16) unsigned char fin_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  wait_for_peers(a.wait);  // (K5) the lists of the other data shards land through NVLink
  int q = blockIdx.x * FIN_WARPS + warp;
  if (a.query_rows != nullptr) {  // a subset of the rows, possibly of a length only the device knows
    int count = a.count;
    if (a.count_dev != nullptr) count = *a.count_dev < count ? *a.count_dev : count;
    if (q >= count) return;
    q = a.query_rows[q];
  }
  if (q >= a.q || q < 0) return;
  int kq = a.k[q];
  kq = kq < 0 ? 0 : (kq > a.kmax ? a.kmax : kq);
  const long long qoff = static_cast<long long>(q) * a.kmax;

  const double *m_dist;
  const int *m_id, *m_label;
  if (a.sources == 1) {
    m_dist = a.dist + qoff;
    m_id = a.id + qoff;
    m_label = a.label + qoff;
  } else {
    double *f_dist;
    int *f_id, *f_label;
    if (a.kmax <= FIN_SMEM_K) {
      const int kpad = (a.kmax + 1) & ~1;
      f_dist = reinterpret_cast<double *>(fin_smem) + static_cast<size_t>(warp) * kpad;
      f_id = reinterpret_cast<int *>(fin_smem + sizeof(double) * FIN_WARPS * kpad) + static_cast<size_t>(warp) * 2 * kpad;
      f_label = f_id + kpad;
    } else {
      unsigned char *w = static_cast<unsigned char *>(a.workspace);
      f_dist = reinterpret_cast<double *>(w) + qoff;
      f_id = reinterpret_cast<int *>(w + sizeof(double) * static_cast<size_t>(a.q) * a.kmax) + qoff;
      f_label = f_id + static_cast<size_t>(a.q) * a.kmax;
    }
    for (int i = lane; i < kq; i += 32) { f_dist[i] = 1e18; f_id[i] = -1; f_label[i] = -1; }
    __syncwarp();
    for (int s = 0; s < a.sources; ++s) {
      const long long soff = s * a.source_stride + qoff;
      const int len = real_entries(a.id + soff, kq);
      for (int p = lane; p < len; p += 32) {
        const double d = a.dist[soff + p];
        const int id = a.id[soff + p];
        int rank = p;
        for (int t = 0; t < a.sources && rank < kq; ++t) {
          if (t == s) continue;
          const long long toff = t * a.source_stride + qoff;
          rank += entries_before(a.dist + toff, a.id + toff, real_entries(a.id + toff, kq), d, id);
        }
        if (rank < kq) { f_dist[rank] = d; f_id[rank] = id; f_label[rank] = a.label[soff + p]; }
      }
    }
    __syncwarp();
    m_dist = f_dist; m_id = f_id; m_label = f_label;
  }

  // majority vote, ties to the larger label, -1 when nothing voted
  unsigned long long best = 0;
  for (int i = lane; i < kq; i += 32) {
    if (m_id[i] < 0) continue;
    const int lab = m_label[i];
    int count = 0;
    for (int j = 0; j < kq; ++j) count += (m_id[j] >= 0 && m_label[j] == lab) ? 1 : 0;
    const unsigned long long key =
        (static_cast<unsigned long long>(count) << 32) | (static_cast<uint32_t>(lab) ^ 0x80000000u);
    best = key > best ? key : best;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    best = other > best ? other : best;
  }
  const int label = best == 0 ? -1 : static_cast<int>(static_cast<uint32_t>(best & 0xFFFFFFFFu) ^ 0x80000000u);

  // checksum: a sequential fold, 32 ids fetched at a time
  unsigned long long h = 1469598103934665603ULL;
  h ^= static_cast<unsigned long long>(static_cast<long long>(label));
  h *= 1099511628211ULL;
  for (int i0 = 0; i0 < kq; i0 += 32) {
    const int mine = i0 + lane < kq ? m_id[i0 + lane] : 0;
    const int cnt = kq - i0 < 32 ? kq - i0 : 32;
    for (int j = 0; j < cnt; ++j) {
      const int id = __shfl_sync(0xFFFFFFFFu, mine, j);
      h ^= static_cast<unsigned long long>(static_cast<long long>(id + 1));
      h *= 1099511628211ULL;
    }
  }
  if (lane == 0) {
    a.out_label[q] = label;
    a.out_checksum[q] = h;
  }
  if (a.out_id != nullptr)
    for (int i = lane; i < a.kmax; i += 32) a.out_id[qoff + i] = i < kq ? m_id[i] : -1;
  if (a.out_dist != nullptr)
    for (int i = lane; i < a.kmax; i += 32) a.out_dist[qoff + i] = i < kq ? m_dist[i] : 1e18;
}

}  // namespace

size_t finalize_workspace_bytes(int sources, int q, int kmax) {
  if (sources <= 1 || kmax <= FIN_SMEM_K) return 0;
  return static_cast<size_t>(q) * kmax * (sizeof(double) + 2 * sizeof(int));
}

int launch_finalize(const FinalizeArgs &args, cudaStream_t s) {
  if (args.q <= 0) return 0;
  DMLP_REQUIRE(args.sources >= 1 && args.kmax >= 1, "finalize: need at least one source and kmax >= 1");
  const size_t need = finalize_workspace_bytes(args.sources, args.q, args.kmax);
  DMLP_REQUIRE(need == 0 || (args.workspace != nullptr && args.workspace_bytes >= need),
               "finalize: merging lists longer than 2048 needs finalize_workspace_bytes() of scratch");
  size_t smem = 0;
  if (args.sources > 1 && args.kmax <= FIN_SMEM_K) {
    const size_t kpad = (static_cast<size_t>(args.kmax) + 1) & ~size_t(1);
    smem = (sizeof(double) + 2 * sizeof(int)) * FIN_WARPS * kpad;
  }
  if (smem > 48 * 1024)
    DMLP_CUDA_OK(cudaFuncSetAttribute(finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  const int rows = args.query_rows != nullptr ? args.count : args.q;
  if (rows <= 0) return 0;
  const int blocks = (rows + FIN_WARPS - 1) / FIN_WARPS;
  finalize_kernel<<<blocks, FIN_WARPS * 32, smem, s>>>(args);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
