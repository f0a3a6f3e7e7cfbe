// K2 exact_rerank: certify the tensor-core filter and restore the reference's
// exact fp64 ordering.
//
// The reference ranks by squared distances accumulated in fp64, strictly in
// index order, without FMA (reference: engine.cpp:12-18); its reported order is
// ascending distance with ties to the larger id (engine.cpp:334-338).  The filter
// GEMM only yields approximate scores, so three kernels run per call:
//
//   select  (one CTA per query) merges the per-chunk key lists, finds the k-th
//           smallest score and derives a threshold T such that every point of the
//           exact top-k must score <= T -- from the measured rounding-error norms
//           of the packed rows (pack.cu) and a bound on the fp32 tensor-core
//           accumulation.  It proves completeness (every full chunk list must have
//           cut off above T: dropped rows score >= the cut-off) or flags the query,
//           and appends the survivors to a dense (query, row) pair list.
//   exact   (one warp per 32 pairs, one lane per pair) recomputes the distance in fp64
//           with the reference's rounding sequence (__dsub_rn/__dmul_rn/__dadd_rn,
//           one sequential chain per lane).  The pass is bound by fp64 instruction
//           issue, so warps are packed full from the dense pair list; rows are
//           streamed in coalesced 256-byte segments through a per-warp cp.async
//           ring, with no block-wide barrier.
//   emit    (one warp per query) checks every recomputed pair against its predicted
//           error, ranks by (distance asc, id desc) and writes the k best as
//           (distance, id, label) -- locally, or (K5) straight into the receive
//           buffer of the rank that owns the query.
// Flagged queries are re-run by the caller on the exact path (exact.cu).
#include <limits>

#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

namespace {

constexpr int RR_THREADS = 128;
// survivors recomputed per query, at most: 256 for the usual small k, about 2 k beyond (k <= 1024:
// the certified survivor set is ~1.6 k on average and its relative spread shrinks as k grows)
inline int wmax_for(int kmax) { return kmax <= 96 ? 256 : (2 * kmax + 64 + 31) / 32 * 32; }

// pair-list capacity per query on average: survivors are about 1.6 k
inline long long pairs_per_query(int kmax) {
  long long v = 2LL * kmax + 16;
  const long long w = wmax_for(kmax);
  return v < 32 ? 32 : (v > w ? w : v);
}

enum : int { kFlagCutoff = 1, kFlagOverflow = 2, kFlagModel = 4 };

struct QMeta {
  double delta, E, qn2, s2;
};

struct Scratch {
  unsigned *counter;
  int *q_base, *q_w;
  QMeta *meta;
  int *pair_q, *pair_id;
  float *pair_score;
  double *pair_dist;
  long long capacity;
};

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

Scratch carve(void *ws, int q, int kmax) {
  Scratch s;
  unsigned char *p = static_cast<unsigned char *>(ws);
  const long long cap = static_cast<long long>(q) * pairs_per_query(kmax) + wmax_for(kmax);
  s.capacity = cap;
  s.counter = reinterpret_cast<unsigned *>(p); p += 256;
  s.q_base = reinterpret_cast<int *>(p); p += up256(sizeof(int) * q);
  s.q_w = reinterpret_cast<int *>(p); p += up256(sizeof(int) * q);
  s.meta = reinterpret_cast<QMeta *>(p); p += up256(sizeof(QMeta) * q);
  s.pair_q = reinterpret_cast<int *>(p); p += up256(sizeof(int) * cap);
  s.pair_id = reinterpret_cast<int *>(p); p += up256(sizeof(int) * cap);
  s.pair_score = reinterpret_cast<float *>(p); p += up256(sizeof(float) * cap);
  s.pair_dist = reinterpret_cast<double *>(p); p += up256(sizeof(double) * cap);
  return s;
}

// ------------------------------------------------------------------ select
__global__ void __launch_bounds__(RR_THREADS) rerank_select_kernel(const RerankArgs a, const Scratch sc, int lpad,
                                                                    int wmax) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  unsigned long long *keys = reinterpret_cast<unsigned long long *>(rr_smem);
  __shared__ int s_nvalid, s_w, s_flag, s_base;
  __shared__ double s_T;

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
    s_T = T;
    QMeta m;
    m.delta = delta; m.E = E; m.qn2 = qn2; m.s2 = sp.scale * sp.scale;
    sc.meta[q] = m;
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
  if (tid == 0) {
    int w = s_w;
    if (w > wmax) { s_flag |= kFlagOverflow; w = wmax; }
    long long base = 0;
    if (w > 0) {
      base = atomicAdd(sc.counter, static_cast<unsigned>(w));
      if (base + w > sc.capacity) { s_flag |= kFlagOverflow; w = 0; base = 0; }  // pair list full
    }
    s_w = w;
    s_base = static_cast<int>(base);
    sc.q_base[q] = static_cast<int>(base);
    sc.q_w[q] = w;
    a.flags[q] = s_flag;
    if (a.w_out != nullptr) a.w_out[q] = w;
  }
  __syncthreads();
  const int w = s_w, base = s_base;
  for (int i = tid; i < w; i += RR_THREADS) {
    sc.pair_q[base + i] = q;
    sc.pair_id[base + i] = key_id(keys[i]);
    sc.pair_score[base + i] = key_score(keys[i]);
  }
}

// ------------------------------------------------------------------- exact
// A warp takes 32 consecutive entries of the dense pair list, whatever queries they belong
// to (full warps); lane r runs pair r's sequential chain.  The warp streams 32-column chunks
// of the 32 data rows and of the lanes' query rows (duplicates are L2 hits) through a private
// two-stage cp.async ring: every global access is a coalesced 256-byte row segment and there
// is no block-wide barrier.
// Measured on B200 (headline shape, 1.03 M pairs x 8192 columns, profiles/ncu_final_kernels_v2_r1.txt):
// 57 GB from HBM in 23.7 M SM cycles = 13.0 ms at 1.83 GHz (4.4 TB/s, 0.67 of the measured copy
// rate), 16-17 ms at the ~1.3 GHz a power-capped step runs at; fp64 pipe 11 % busy, 6 warps per
// SM.  Tried and slower: 16-column chunks with 2, 3 or 4 stages (17.7-18.3 ms), 32 columns x 3
// stages (25.7 ms, 4 warps per SM), one cp.async.bulk per lane and row segment (50 ms: the
// copy engine is per-warp uniform and 256-byte copies are far below its efficient size),
// chains spread over 4 warps, per-thread global loads.
constexpr int XC = 32;             // columns per chunk
constexpr int XSTR = XC + 2;       // row stride in doubles (16-byte aligned rows)
constexpr int XROWS = 64;          // 32 data rows + 32 query rows
constexpr int XSTAGE = XROWS * XSTR;
constexpr int X_WARPS = 2;
constexpr int X_SMEM = X_WARPS * 2 * XSTAGE * 8;

template <int BYTES>
__device__ __forceinline__ void cp_async_zfill(void *dst, const void *src, bool valid) {
  const int sz = valid ? BYTES : 0;  // src-size 0 zero-fills the destination
  if (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}

template <bool VEC2>
__global__ void __launch_bounds__(X_WARPS * 32) rerank_exact_kernel(const RerankArgs a, const Scratch sc) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned count = *sc.counter;
  const long long total = count < sc.capacity ? static_cast<long long>(count) : sc.capacity;
  const long long p0 = (static_cast<long long>(blockIdx.x) * X_WARPS + warp) * 32;
  if (p0 >= total) return;
  const int cnt = (total - p0) < 32 ? static_cast<int>(total - p0) : 32;
  double *ring = reinterpret_cast<double *>(rr_smem) + static_cast<size_t>(warp) * 2 * XSTAGE;
  // this lane's pair: element offsets of its data row and of its query row
  const long long my_x = lane < cnt ? static_cast<long long>(sc.pair_id[p0 + lane]) * a.a : 0;
  const long long my_q = lane < cnt ? static_cast<long long>(sc.pair_q[p0 + lane]) * a.a : 0;
  const int nchunks = (a.a + XC - 1) / XC;

  // stage rows 0..31: data rows of lanes 0..31; rows 32..63: their query rows
  auto issue = [&](int chunk) {
    if (chunk < nchunks) {
      double *stage = ring + (chunk & 1) * XSTAGE;
      const int col0 = chunk * XC;
      if (VEC2) {
        // 16 lanes x 16 B cover one 32-column row segment; two rows per instruction
        const int half = lane >> 4, seg = lane & 15;
        const int col = col0 + seg * 2;
        const bool in = col + 2 <= a.a;
        for (int r0 = 0; r0 < cnt; r0 += 2) {
          const int r = r0 + half;
          const long long xo = __shfl_sync(0xFFFFFFFFu, my_x, r < cnt ? r : 0);
          const long long qo = __shfl_sync(0xFFFFFFFFu, my_q, r < cnt ? r : 0);
          if (r < cnt) {
            cp_async_zfill<16>(stage + r * XSTR + seg * 2, a.x_attrs + xo + (in ? col : 0), in);
            cp_async_zfill<16>(stage + (32 + r) * XSTR + seg * 2, a.q_attrs + qo + (in ? col : 0), in);
          }
        }
      } else {
        const int col = col0 + lane;
        const bool in = col < a.a;
        for (int r = 0; r < cnt; ++r) {
          const long long xo = __shfl_sync(0xFFFFFFFFu, my_x, r);
          const long long qo = __shfl_sync(0xFFFFFFFFu, my_q, r);
          cp_async_zfill<8>(stage + r * XSTR + lane, a.x_attrs + xo + (in ? col : 0), in);
          cp_async_zfill<8>(stage + (32 + r) * XSTR + lane, a.q_attrs + qo + (in ? col : 0), in);
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");  // always commit: keeps the group count uniform
  };

  double acc = 0.0;
  issue(0);
  for (int chunk = 0; chunk < nchunks; ++chunk) {
    issue(chunk + 1);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    if (lane < cnt) {
      const double *stage = ring + (chunk & 1) * XSTAGE;
      const double *xs = stage + lane * XSTR;
      const double *qs = stage + (32 + lane) * XSTR;
#pragma unroll 8
      for (int i = 0; i < XC; ++i) {
        const double d = __dsub_rn(qs[i], xs[i]);
        acc = __dadd_rn(acc, __dmul_rn(d, d));  // zero-filled tail adds +0.0 exactly
      }
    }
    __syncwarp();
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (lane < cnt) sc.pair_dist[p0 + lane] = acc;
}

// Variant that stages every distinct query row once.  The dense pair list keeps a query's
// pairs together, so the 32 pairs of a warp belong to a few queries only (about two at
// k ~ 10); the ring then needs 32 data rows + XQ query rows per stage instead of 64 rows:
// less shared memory per warp (10 instead of 6 resident warps per SM), fewer staging
// instructions and 40 % less L2 -> SM traffic.  A warp whose pairs span more than XQ queries
// (tiny k) handles them XQ queries at a time.  Measured on the headline shape inside a step:
// 17.3 -> 11.4 ms (profiles/sweep_rerank_dedup_r1.log).
constexpr int XQ = 8;
constexpr int XD_ROWS = 32 + XQ;
constexpr int XD_STAGE = XD_ROWS * XSTR;
constexpr int XD_SMEM = X_WARPS * 2 * XD_STAGE * 8;
// XI32: the data rows are resident in decimal6 form (int32 i, value i / 1e6): half the bytes of
// the pass, which is bound by exactly those bytes.  A stage then holds 32 int rows (stride XSTRI)
// followed by the XQ query rows (fp64); the value is rebuilt with dec6_to_double (3 fp64 operations).
constexpr int XSTRI = XC + 4;                                  // ints; rows stay 16-byte aligned (and LDS.128 by 8 lanes is conflict-free)
constexpr int XI_DATA_BYTES = 32 * XSTRI * 4;                  // 4608
constexpr int XI_STAGE_BYTES = XI_DATA_BYTES + XQ * XSTR * 8;  // + 2176
constexpr int XI_SMEM = X_WARPS * 2 * XI_STAGE_BYTES;
// XMODE 0: fp64 rows.  1: decimal6 rows, staged two rows per warp instruction (8-byte pieces, any even a).
// 2: decimal6 rows with a % 4 == 0: four rows per warp instruction (16-byte pieces) and the lanes read
// their row back four values at a time (LDS.128, conflict-free at this row stride) -- fewer instructions,
// which is what bounds the pass once the bytes are halved.  (Letting every lane copy its own row segment
// -- no shuffles at all -- was 6x slower: 32 scattered 16-byte requests per instruction.)

template <int XMODE>
__global__ void __launch_bounds__(X_WARPS * 32) rerank_exact_dedup_kernel(const RerankArgs a, const Scratch sc) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned count = *sc.counter;
  const long long total = count < sc.capacity ? static_cast<long long>(count) : sc.capacity;
  const long long p0 = (static_cast<long long>(blockIdx.x) * X_WARPS + warp) * 32;
  if (p0 >= total) return;
  const int cnt = (total - p0) < 32 ? static_cast<int>(total - p0) : 32;
  constexpr bool XI32 = XMODE != 0;
  constexpr int STAGE_BYTES = XI32 ? XI_STAGE_BYTES : XD_STAGE * 8;
  unsigned char *ring = rr_smem + static_cast<size_t>(warp) * 2 * STAGE_BYTES;
  const bool have = lane < cnt;
  const int my_qi = have ? sc.pair_q[p0 + lane] : -1;
  const long long my_x = have ? static_cast<long long>(sc.pair_id[p0 + lane]) * a.a : 0;
  // runs of equal query index -> slot number of this lane's query, and the lane that starts each run
  const int prev_qi = __shfl_up_sync(0xFFFFFFFFu, my_qi, 1);
  const unsigned starts = __ballot_sync(0xFFFFFFFFu, have && (lane == 0 || my_qi != prev_qi));
  const int slot = __popc(starts & (0xFFFFFFFFu >> (31 - lane))) - 1;
  const int distinct = __popc(starts);
  // lane s (< distinct) holds the row offset of the query of slot s
  const int start_lane = lane < distinct ? static_cast<int>(__fns(starts, 0, lane + 1)) : 0;
  const long long slot_q = static_cast<long long>(__shfl_sync(0xFFFFFFFFu, my_qi, start_lane)) * a.a;
  const int nchunks = (a.a + XC - 1) / XC;
  const int sub = lane >> 4, seg = lane & 15;  // 16 lanes cover one 32-column row segment

  double acc = 0.0;
  for (int base = 0; base < distinct; base += XQ) {
    const bool active = have && slot >= base && slot < base + XQ;
    const unsigned amask = __ballot_sync(0xFFFFFFFFu, active);
    const int nq = min(XQ, distinct - base);
    auto issue = [&](int chunk) {
      if (chunk < nchunks) {
        unsigned char *stage = ring + (chunk & 1) * STAGE_BYTES;
        double *qstage = reinterpret_cast<double *>(stage + (XI32 ? XI_DATA_BYTES : 32 * XSTR * 8));
        const int col = chunk * XC + seg * 2;
        const bool in = col + 2 <= a.a;
        if (XMODE == 2) {
          // 8 lanes x 16 B cover one 32-column row segment: four rows per warp instruction, coalesced
          const int sub4 = lane >> 3, seg4 = lane & 7;
          const int col4 = chunk * XC + seg4 * 4;
          const bool in4 = col4 + 4 <= a.a;
          for (int r0 = 0; r0 < cnt; r0 += 4) {
            const int r = r0 + sub4;
            const long long xo = __shfl_sync(0xFFFFFFFFu, my_x, r < cnt ? r : 0);
            if (r < cnt && ((amask >> r) & 1u))
              cp_async_zfill<16>(reinterpret_cast<int *>(stage) + r * XSTRI + seg4 * 4, a.x_attrs_i32 + xo + (in4 ? col4 : 0), in4);
          }
        } else
        for (int r0 = 0; r0 < cnt; r0 += 2) {  // data rows of the lanes handled in this pass
          const int r = r0 + sub;
          const long long xo = __shfl_sync(0xFFFFFFFFu, my_x, r < cnt ? r : 0);
          if (r < cnt && ((amask >> r) & 1u)) {
            if (XI32)
              cp_async_zfill<8>(reinterpret_cast<int *>(stage) + r * XSTRI + seg * 2, a.x_attrs_i32 + xo + (in ? col : 0), in);
            else
              cp_async_zfill<16>(reinterpret_cast<double *>(stage) + r * XSTR + seg * 2, a.x_attrs + xo + (in ? col : 0), in);
          }
        }
        for (int s0 = 0; s0 < nq; s0 += 2) {   // one copy of each query row
          const int sl = s0 + sub;
          const long long qo = __shfl_sync(0xFFFFFFFFu, slot_q, base + (sl < nq ? sl : 0));
          if (sl < nq) cp_async_zfill<16>(qstage + sl * XSTR + seg * 2, a.q_attrs + qo + (in ? col : 0), in);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");  // always commit: keeps the group count uniform
    };
    issue(0);
    for (int chunk = 0; chunk < nchunks; ++chunk) {
      issue(chunk + 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncwarp();
      if (active) {
        const unsigned char *stage = ring + (chunk & 1) * STAGE_BYTES;
        const double *qs = reinterpret_cast<const double *>(stage + (XI32 ? XI_DATA_BYTES : 32 * XSTR * 8)) + (slot - base) * XSTR;
        if (XMODE == 2) {
          const int4 *xs4 = reinterpret_cast<const int4 *>(reinterpret_cast<const int *>(stage) + lane * XSTRI);
          const double2 *qs2 = reinterpret_cast<const double2 *>(qs);
#pragma unroll 4
          for (int i = 0; i < XC / 4; ++i) {
            const int4 xv = xs4[i];
            const double2 qa = qs2[2 * i], qb = qs2[2 * i + 1];
            double d = __dsub_rn(qa.x, dec6_to_double(xv.x));
            acc = __dadd_rn(acc, __dmul_rn(d, d));
            d = __dsub_rn(qa.y, dec6_to_double(xv.y));
            acc = __dadd_rn(acc, __dmul_rn(d, d));
            d = __dsub_rn(qb.x, dec6_to_double(xv.z));
            acc = __dadd_rn(acc, __dmul_rn(d, d));
            d = __dsub_rn(qb.y, dec6_to_double(xv.w));
            acc = __dadd_rn(acc, __dmul_rn(d, d));
          }
        } else if (XI32) {
          const int *xs = reinterpret_cast<const int *>(stage) + lane * XSTRI;
#pragma unroll 8
          for (int i = 0; i < XC; ++i) {
            // (a zero-filled tail is the integer 0 -> 0.0 against a query value of 0.0: adds +0.0 exactly)
            const double d = __dsub_rn(qs[i], dec6_to_double(xs[i]));
            acc = __dadd_rn(acc, __dmul_rn(d, d));
          }
        } else {
          const double *xs = reinterpret_cast<const double *>(stage) + lane * XSTR;
#pragma unroll 8
          for (int i = 0; i < XC; ++i) {
            const double d = __dsub_rn(qs[i], xs[i]);
            acc = __dadd_rn(acc, __dmul_rn(d, d));  // zero-filled tail adds +0.0 exactly
          }
        }
      }
      __syncwarp();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
  }
  if (have) sc.pair_dist[p0 + lane] = acc;
}

// -------------------------------------------------------------------- emit
// With `final_label` set (single data shard: the local list is the final one) the warp also
// votes and folds the checksum right here -- the K6 tail (reference: engine.cpp:320-347,
// common.cpp:59-68) without a second kernel or a second pass over the lists.
__device__ __forceinline__ void emit_query(const RerankArgs &a, const Scratch &sc, int q, int lane, double *s_dist,
                                           int *s_id, int *s_fid, int *s_flab) {
  int kq = a.k[q];
  kq = kq < 0 ? 0 : (kq > a.kmax ? a.kmax : kq);
  const int w = sc.q_w[q], base = sc.q_base[q];
  const QMeta m = sc.meta[q];
  const bool fold = a.final_label != nullptr && a.scatter.world <= 1;
  int bad = 0;
  for (int i = lane; i < w; i += 32) {
    const double d = sc.pair_dist[base + i];
    s_dist[i] = d;
    s_id[i] = sc.pair_id[base + i];
    // a-posteriori check of the error model on every recomputed pair
    const double dhat = fmax(m.qn2 + static_cast<double>(sc.pair_score[base + i]), 0.0);
    const double bound = (2.0 * sqrt(dhat) * m.delta + m.delta * m.delta + m.E) * (1.0 + 1e-6) + 1e-9 * dhat;
    if (fabs(d * m.s2 - dhat) > bound) bad = 1;
  }
  if (fold)
    for (int i = lane; i < kq; i += 32) { s_fid[i] = -1; s_flab[i] = -1; }
  bad = __any_sync(0xFFFFFFFFu, bad);
  __syncwarp();

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
  for (int i = lane; i < w; i += 32) {
    const double d = s_dist[i];
    const int id = s_id[i];
    int rank = 0;
    for (int j = 0; j < w; ++j) {
      const double dj = s_dist[j];
      rank += (dj < d || (dj == d && s_id[j] > id)) ? 1 : 0;
    }
    if (rank < kq) {
      const int lab = a.x_labels[id];
      o_dist[obase + rank] = d;
      o_id[obase + rank] = a.id_base + id;
      o_label[obase + rank] = lab;
      if (fold) { s_fid[rank] = a.id_base + id; s_flab[rank] = lab; }
    }
  }
  for (int i = (w < kq ? w : kq) + lane; i < a.kmax; i += 32) {
    o_dist[obase + i] = 1e18;
    o_id[obase + i] = -1;
    o_label[obase + i] = -1;
  }
  if (lane == 0) {
    const int f = a.flags[q] | (bad ? kFlagModel : 0);
    a.flags[q] = f;
    if (f != 0 && a.num_flagged != nullptr) {
      const int slot = atomicAdd(a.num_flagged, 1);
      if (a.flagged_rows != nullptr && slot < a.flagged_cap) a.flagged_rows[slot] = q;
    }
  }
  if (!fold) return;
  __syncwarp();
  // majority vote, ties to the larger label, -1 when nothing voted
  unsigned long long best = 0;
  for (int i = lane; i < kq; i += 32) {
    if (s_fid[i] < 0) continue;
    const int lab = s_flab[i];
    int count = 0;
    for (int j = 0; j < kq; ++j) count += (s_fid[j] >= 0 && s_flab[j] == lab) ? 1 : 0;
    const unsigned long long key =
        (static_cast<unsigned long long>(count) << 32) | (static_cast<uint32_t>(lab) ^ 0x80000000u);
    best = key > best ? key : best;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, best, o);
    best = other > best ? other : best;
  }
  if (lane == 0) {
    const int label = best == 0 ? -1 : static_cast<int>(static_cast<uint32_t>(best & 0xFFFFFFFFu) ^ 0x80000000u);
    unsigned long long h = 1469598103934665603ULL;
    h ^= static_cast<unsigned long long>(static_cast<long long>(label));
    h *= 1099511628211ULL;
    for (int i = 0; i < kq; ++i) {
      h ^= static_cast<unsigned long long>(static_cast<long long>(s_fid[i] + 1));
      h *= 1099511628211ULL;
    }
    a.final_label[q] = label;
    a.final_checksum[q] = h;
  }
}


// shared memory per warp: wmax distances + wmax ids (the survivors), kmax ids + kmax labels (the ranked list)
inline size_t emit_smem_per_warp(int wmax, int kmax) {
  return static_cast<size_t>(wmax) * 12 + (static_cast<size_t>(kmax + 1) / 2 * 2) * 8;
}

__global__ void __launch_bounds__(RR_THREADS) rerank_emit_kernel(const RerankArgs a, const Scratch sc, int wmax) {
  extern __shared__ __align__(16) unsigned char rr_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kpad = (a.kmax + 1) / 2 * 2;
  unsigned char *mine = rr_smem + static_cast<size_t>(warp) * (static_cast<size_t>(wmax) * 12 + static_cast<size_t>(kpad) * 8);
  double *s_dist = reinterpret_cast<double *>(mine);
  int *s_id = reinterpret_cast<int *>(mine + static_cast<size_t>(wmax) * 8);
  int *s_fid = s_id + wmax;
  int *s_flab = s_fid + kpad;
  const int q = blockIdx.x * (RR_THREADS / 32) + warp;
  if (q < a.q) emit_query(a, sc, q, lane, s_dist, s_id, s_fid, s_flab);
  signal_peers_when_grid_done(a.signal);  // (K5) all of this rank's lists have been stored
}

int *g_debug_w = nullptr;

}  // namespace

void set_rerank_debug_w(int *w_out) { g_debug_w = w_out; }

size_t rerank_workspace_bytes(int q, int kmax) {
  if (q < 1) q = 1;
  const size_t cap = static_cast<size_t>(q) * pairs_per_query(kmax) + wmax_for(kmax);
  return 256 + 2 * up256(sizeof(int) * q) + up256(sizeof(QMeta) * q) + 2 * up256(sizeof(int) * cap) +
         up256(sizeof(float) * cap) + up256(sizeof(double) * cap);
}

int launch_rerank(const RerankArgs &args_in, cudaStream_t s) {
  RerankArgs args = args_in;
  if (args.w_out == nullptr) args.w_out = g_debug_w;
  if (args.q <= 0) return 0;
  const int L = args.chunks * args.kp;
  DMLP_REQUIRE(L > 0 && L <= 8192, "rerank: chunks * kp must be in (0, 8192]");
  DMLP_REQUIRE(args.kmax > 0, "rerank: kmax must be positive");
  DMLP_REQUIRE(args.workspace != nullptr && args.workspace_bytes >= rerank_workspace_bytes(args.q, args.kmax),
               "rerank: workspace missing or smaller than rerank_workspace_bytes(q, kmax)");
  DMLP_REQUIRE((reinterpret_cast<uintptr_t>(args.workspace) & 255) == 0, "rerank: workspace must be 256-byte aligned");
  const Scratch sc = carve(args.workspace, args.q, args.kmax);
  DMLP_CUDA_OK(cudaMemsetAsync(sc.counter, 0, sizeof(unsigned), s));
  int lpad = 64;
  while (lpad < L) lpad <<= 1;
  const size_t smem = static_cast<size_t>(lpad) * 8;
  DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  const int wmax = wmax_for(args.kmax);
  rerank_select_kernel<<<args.q, RR_THREADS, smem, s>>>(args, sc, lpad, wmax);
  DMLP_CUDA_OK(cudaGetLastError());
  // one warp per 32 pairs; sized for the worst case, surplus warps exit at once
  const unsigned blocks = static_cast<unsigned>((sc.capacity + 32 * X_WARPS - 1) / (32 * X_WARPS));
  const bool vec2 = (args.a % 2 == 0) && (reinterpret_cast<uintptr_t>(args.q_attrs) % 16 == 0) &&
                    (reinterpret_cast<uintptr_t>(args.x_attrs) % 16 == 0);
  static const bool dedup = []() {
    const char *e = getenv("DMLP_RERANK_DEDUP");  // 0: stage a query row per pair (the first version)
    return e == nullptr || atoi(e) != 0;
  }();
  DMLP_REQUIRE(args.x_attrs_i32 == nullptr ||
                   ((args.a % 2 == 0) && reinterpret_cast<uintptr_t>(args.x_attrs_i32) % 8 == 0 &&
                    reinterpret_cast<uintptr_t>(args.q_attrs) % 16 == 0),
               "rerank: decimal6 rows need an even row length and aligned buffers");
  if (args.x_attrs_i32 != nullptr && args.a % 4 == 0 && reinterpret_cast<uintptr_t>(args.x_attrs_i32) % 16 == 0) {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_exact_dedup_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, XI_SMEM));
    rerank_exact_dedup_kernel<2><<<blocks, X_WARPS * 32, XI_SMEM, s>>>(args, sc);
  } else if (args.x_attrs_i32 != nullptr) {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_exact_dedup_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, XI_SMEM));
    rerank_exact_dedup_kernel<1><<<blocks, X_WARPS * 32, XI_SMEM, s>>>(args, sc);
  } else if (vec2 && dedup) {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_exact_dedup_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, XD_SMEM));
    rerank_exact_dedup_kernel<0><<<blocks, X_WARPS * 32, XD_SMEM, s>>>(args, sc);
  } else if (vec2) {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_exact_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, X_SMEM));
    rerank_exact_kernel<true><<<blocks, X_WARPS * 32, X_SMEM, s>>>(args, sc);
  } else {
    DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_exact_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, X_SMEM));
    rerank_exact_kernel<false><<<blocks, X_WARPS * 32, X_SMEM, s>>>(args, sc);
  }
  DMLP_CUDA_OK(cudaGetLastError());
  const size_t emit_smem = (RR_THREADS / 32) * emit_smem_per_warp(wmax, args.kmax);
  DMLP_REQUIRE(emit_smem <= 227 * 1024, "rerank: kmax too large for the emit kernel's shared memory");
  DMLP_CUDA_OK(cudaFuncSetAttribute(rerank_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(emit_smem)));
  rerank_emit_kernel<<<(args.q + RR_THREADS / 32 - 1) / (RR_THREADS / 32), RR_THREADS, emit_smem, s>>>(args, sc, wmax);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
