// K2 exact_rerank: certify the tensor-core filter and restore the reference's
// exact fp64 ordering.
//
// The reference ranks by squared distances accumulated in fp64, strictly in
// index order, without FMA (reference: engine.cpp:12-18); its reported order is
// ascending distance with ties to the larger id (engine.cpp:334-338).  The filter
// GEMM only yields approximate scores, so for each query this kernel
//   1. merges the per-chunk key lists and finds the k-th smallest score,
//   2. derives a score threshold T such that every point of the exact top-k must
//      have score <= T, from the measured rounding-error norms of the packed
//      rows (pack.cu) and a bound on the fp32 tensor-core accumulation,
//   3. proves completeness: every chunk whose list is full must have cut off
//      above T (dropped rows score >= the cut-off); otherwise the query is
//      flagged and the caller re-runs it on the exact path (exact.cu),
//   4. recomputes the survivors in fp64 with the reference's rounding sequence
//      (__dsub_rn/__dmul_rn/__dadd_rn, one sequential chain per candidate),
//      checks each against its predicted error, ranks them and writes the k
//      best as (distance, id, label).
#include <limits>

#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

namespace {

constexpr int RR_THREADS = 128;
constexpr int WMAX = 64;       // survivors recomputed per query
constexpr int CH = 32;         // columns staged per step
constexpr int RSTRIDE = CH + 2;  // doubles; keeps rows 16-byte aligned

enum : int { kFlagCutoff = 1, kFlagOverflow = 2, kFlagModel = 4 };

template <int BYTES>
__device__ __forceinline__ void cp_async(void *dst, const void *src, bool valid) {
  const int sz = valid ? BYTES : 0;  // src-size 0 zero-fills the destination
  if (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int VEC>  // doubles per cp.async: 2 (16 B) or 1 (8 B)
__global__ void __launch_bounds__(RR_THREADS) rerank_kernel(const RerankArgs a, int lpad) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  unsigned long long *keys = reinterpret_cast<unsigned long long *>(rr_smem);
  double *xbuf = reinterpret_cast<double *>(rr_smem + static_cast<size_t>(lpad) * 8);  // [2][WMAX+1][RSTRIDE]
  __shared__ double s_dist[WMAX];
  __shared__ int s_id[WMAX];
  __shared__ int s_nvalid, s_w, s_flag;
  __shared__ double s_T, s_delta, s_E, s_qn2, s_s2;

  const int tid = threadIdx.x;
  const int q = blockIdx.x;
  const int L = a.chunks * a.kp;
  const unsigned long long *my_cand = a.cand + static_cast<long long>(q) * L;
  int kq = a.k[q];
  kq = kq < 0 ? 0 : (kq > a.kmax ? a.kmax : kq);

  for (int i = tid; i < lpad; i += RR_THREADS) keys[i] = i < L ? my_cand[i] : kKeyInvalid;
  if (tid == 0) { s_nvalid = 0; s_w = 0; s_flag = 0; }
  __syncthreads();

  // block-wide bitonic sort, ascending
  for (int kk = 2; kk <= lpad; kk <<= 1)
    for (int j = kk >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < lpad; i += RR_THREADS) {
        const int p = i ^ j;
        if (p > i) {
          unsigned long long x = keys[i], y = keys[p];
          const bool asc = (i & kk) == 0;
          if ((x > y) == asc) { keys[i] = y; keys[p] = x; }
        }
      }
      __syncthreads();
    }

  for (int i = tid; i < lpad; i += RR_THREADS)
    if (keys[i] != kKeyInvalid && (i + 1 == lpad || keys[i + 1] == kKeyInvalid)) s_nvalid = i + 1;
  __syncthreads();
  const int nvalid = s_nvalid;
  const int kk_eff = kq < nvalid ? kq : nvalid;

  if (tid == 0) {
    const ScaleParams sp = derive_scale(a.range->neg_lo, a.range->hi);
    const double qn2 = static_cast<double>(a.q_norm2[q]);
    const double delta = static_cast<double>(a.q_errn[q]) + static_cast<double>(a.x_maxima->max_err);
    const double xmax = static_cast<double>(a.x_maxima->max_norm);
    const double qn = sqrt(qn2);
    // fp32 accumulation in the tensor core: one rounding (towards zero at worst)
    // per 16-deep MMA step on a partial sum bounded by ||q^|| ||x^||, doubled as
    // a safety margin; plus fp32 rounding of norms and of the fused score.
    const double e_acc = 2.0 * (a.a_pad / 16.0 + 2.0) * 2.384185791015625e-07 * qn * xmax;
    const double e_f32 = 16.0 * 5.9604644775390625e-08 * (qn2 + xmax * xmax + 2.0 * qn * xmax);
    const double E = e_acc + e_f32;
    double T = kq == 0 ? -std::numeric_limits<double>::infinity() : std::numeric_limits<double>::infinity();
    if (kk_eff > 0 && kk_eff == kq) {
      const double sk = static_cast<double>(key_score(keys[kk_eff - 1]));
      const double dk = fmax(qn2 + sk, 0.0);
      const double U = dk + 2.0 * sqrt(dk) * delta + delta * delta + E;
      const double root = delta + sqrt(2.0 * delta * delta + E + U);
      T = root * root * (1.0 + 1e-6) - qn2;
    }
    s_T = T; s_delta = delta; s_E = E; s_qn2 = qn2;
    s_s2 = sp.scale * sp.scale;
  }
  __syncthreads();
  const double T = s_T;

  // survivors form a prefix of the sorted keys
  for (int i = tid; i < nvalid; i += RR_THREADS) {
    const bool in = static_cast<double>(key_score(keys[i])) <= T;
    const bool next_in = (i + 1 < nvalid) && static_cast<double>(key_score(keys[i + 1])) <= T;
    if (in && !next_in) s_w = i + 1;
  }
  // completeness: a full chunk list must have cut off strictly above T
  for (int c = tid; c < a.chunks; c += RR_THREADS) {
    const unsigned long long last = my_cand[c * a.kp + a.kp - 1];
    if (last != kKeyInvalid && !(static_cast<double>(key_score(last)) > T)) atomicOr(&s_flag, kFlagCutoff);
  }
  __syncthreads();
  int w = s_w;
  if (w > WMAX) {
    if (tid == 0) atomicOr(&s_flag, kFlagOverflow);
    w = WMAX;
  }

  // ---- exact fp64 distances for the w survivors -----------------------------
  const double *qrow = a.q_attrs + static_cast<long long>(q) * a.a;
  const int my_id = tid < w ? key_id(keys[tid]) : 0;
  const int nsteps = (a.a + CH - 1) / CH;
  constexpr int SEG = CH / VEC;  // copies per row per step
  auto issue = [&](int step, int buf) {
    const int col0 = step * CH;
    double *dst_base = xbuf + static_cast<size_t>(buf) * (WMAX + 1) * RSTRIDE;
    for (int sidx = tid; sidx < (w + 1) * SEG; sidx += RR_THREADS) {
      const int row = sidx / SEG, off = (sidx - row * SEG) * VEC;
      const double *src_row = row == 0 ? qrow : a.x_attrs + static_cast<long long>(key_id(keys[row - 1])) * a.a;
      const int col = col0 + off;
      const bool valid = col + VEC <= a.a;
      const double *src = src_row + (valid ? col : 0);
      double *dst = dst_base + row * RSTRIDE + off;
      if (VEC == 2 && !valid && col < a.a) {
        // odd tail: one real element, then zero
        dst[0] = src_row[col];
        dst[1] = 0.0;
      } else {
        cp_async<VEC * 8>(dst, src, valid);
      }
    }
    cp_async_commit();
  };

  double acc = 0.0;
  if (w > 0) {
    issue(0, 0);
    for (int step = 0; step < nsteps; ++step) {
      const int buf = step & 1;
      if (step + 1 < nsteps) {
        issue(step + 1, buf ^ 1);
        cp_async_wait<1>();
      } else {
        cp_async_wait<0>();
      }
      __syncthreads();
      if (tid < w) {
        const double *qs = xbuf + static_cast<size_t>(buf) * (WMAX + 1) * RSTRIDE;
        const double *xs = qs + (tid + 1) * RSTRIDE;
#pragma unroll 8
        for (int i = 0; i < CH; ++i) {
          const double d = __dsub_rn(qs[i], xs[i]);
          acc = __dadd_rn(acc, __dmul_rn(d, d));  // zero-filled tail adds +0.0 exactly
        }
      }
      __syncthreads();
    }
  }
  if (tid < w) {
    s_dist[tid] = acc;
    s_id[tid] = my_id;
    // a-posteriori check of the error model on every recomputed pair
    const double dhat = fmax(s_qn2 + static_cast<double>(key_score(keys[tid])), 0.0);
    const double bound = (2.0 * sqrt(dhat) * s_delta + s_delta * s_delta + s_E) * (1.0 + 1e-6) + 1e-9 * dhat;
    if (fabs(acc * s_s2 - dhat) > bound) atomicOr(&s_flag, kFlagModel);
  }
  __syncthreads();

  // rank by (distance asc, id desc) and emit the first kq -- locally, or (K5) straight
  // into the receive buffer of the rank that owns this query, over NVLink
  long long obase = static_cast<long long>(q) * a.kmax;
  double *o_dist = a.out_dist;
  int *o_id = a.out_id, *o_label = a.out_label;
  if (a.scatter.world > 1) {
    const int owner = q / a.scatter.slice_rows;
    const int ql = q - owner * a.scatter.slice_rows;
    obase = (static_cast<long long>(a.scatter.rank) * a.scatter.slice_rows + ql) * a.kmax;
    o_dist = reinterpret_cast<double *>(a.scatter.peer_base[owner] + a.scatter.off_dist);
    o_id = reinterpret_cast<int *>(a.scatter.peer_base[owner] + a.scatter.off_id);
    o_label = reinterpret_cast<int *>(a.scatter.peer_base[owner] + a.scatter.off_label);
  }
  if (tid < w) {
    const double d = s_dist[tid];
    const int id = s_id[tid];
    int rank = 0;
    for (int j = 0; j < w; ++j) {
      const double dj = s_dist[j];
      rank += (dj < d || (dj == d && s_id[j] > id)) ? 1 : 0;
    }
    if (rank < kq) {
      o_dist[obase + rank] = d;
      o_id[obase + rank] = a.id_base + id;
      o_label[obase + rank] = a.x_labels[id];
    }
  }
  for (int i = (w < kq ? w : kq) + tid; i < a.kmax; i += RR_THREADS) {
    o_dist[obase + i] = 1e18;
    o_id[obase + i] = -1;
    o_label[obase + i] = -1;
  }
  if (tid == 0) {
    a.flags[q] = s_flag;
    if (s_flag != 0 && a.num_flagged != nullptr) atomicAdd(a.num_flagged, 1);
  }
}

}  // namespace

int launch_rerank(const RerankArgs &args, cudaStream_t s) {
  if (args.q <= 0) return 0;
  const int L = args.chunks * args.kp;
  DMLP_REQUIRE(L > 0 && L <= 8192, "rerank: chunks * kp must be in (0, 8192]");
  DMLP_REQUIRE(args.kmax > 0, "rerank: kmax must be positive");
  int lpad = 64;
  while (lpad < L) lpad <<= 1;
  const size_t smem = static_cast<size_t>(lpad) * 8 + sizeof(double) * 2 * (WMAX + 1) * RSTRIDE;
  const bool vec2 = (args.a % 2 == 0) && (reinterpret_cast<uintptr_t>(args.q_attrs) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(args.x_attrs) % 16 == 0);
  if (vec2) {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    rerank_kernel<2><<<args.q, RR_THREADS, smem, s>>>(args, lpad);
  } else {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    rerank_kernel<1><<<args.q, RR_THREADS, smem, s>>>(args, lpad);
  }
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
