// K1 + K3: tcgen05 distance GEMM with the per-query top-k selection fused into
// the tensor-memory epilogue.
//
// Replaces the reference's hot loop -- `q x n` calls of computeDistance followed
// by std::nth_element per query (reference: engine.cpp:235-256).  Here the
// cross term of ||q - x||^2 = ||q||^2 + ||x||^2 - 2 q.x is a K-major fp16 GEMM
// on the 5th-generation tensor cores; the full q x n score matrix never exists
// in memory.  Per CTA:
//
//   warp 0      TMA producer: 128x64 query tile + 256x64 data tile per stage,
//               128-byte swizzle, mbarrier full/empty ring
//   warp 1      MMA issuer: one thread issues tcgen05.mma 128x256x16 (x4 per
//               stage) into one of two 256-column accumulators in tensor memory
//   warps 2-5   epilogue: tcgen05.ld 32 lanes x 32 columns, score =
//               ||x^||^2 - 2 acc, threshold test against the row's running
//               k'-th best, append survivors to the row's list in shared
//               memory, warp-cooperative bitonic compaction when a list fills
//
// A CTA sweeps one 128-query block over a contiguous range of data tiles (a
// "chunk") so the running lists and thresholds stay on chip for the whole
// sweep; each (query, chunk) pair emits its `kp` best (score, row) keys sorted
// ascending.  Rows a chunk dropped all score >= the last key it kept, which is
// what rerank.cu needs to certify the filter.
//
// Three kernels share the epilogue: dist_topk_kernel (one CTA per 128 x 256 tile,
// used when there is a single row block), dist_topk_pair_kernel<false> (CTA pair,
// tcgen05 cta_group::2, 256 x 256 per pair and k-step, two accumulator stages so the
// drain of one tile overlaps the MMAs of the next) and dist_topk_pair_kernel<true>
// (the "wide" tile for long rows: one query tile against two data tiles per k-block,
// 256 x 512 per pair, both accumulator stages filled side by side).  The pair kernels
// also carry the communication warp of the fused all-gathers (K4, data rows) and the
// per-tile re-alignment of the pairs that sweep the same chunk.
#include <atomic>
#include <cstdlib>
#include <limits>

#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

namespace {

constexpr int BM = 128;   // queries per tile (UMMA M)
constexpr int BN = 256;   // datapoints per tile (UMMA N)
constexpr int BK = 64;    // fp16 elements per stage along K (= 128 B swizzle span)
constexpr int STAGES = 3;
constexpr int A_BYTES = BM * BK * 2;
constexpr int B_BYTES = BN * BK * 2;
constexpr int KP = 32;    // keys kept per (query, chunk)
constexpr int CAP = 64;   // list capacity between compactions
constexpr int THREADS = 192;
constexpr int EPI_THREADS = 128;


// Re-alignment waits that gave up (see dist_topk_sync_timeouts): results stay correct, the
// sweep only loses its L2 sharing, but a non-zero count means launches overlapped on one
// slice of the counters or a CTA pair never arrived.
__device__ unsigned g_sync_timeouts = 0;

struct Item {
  int m;
  int chunk;
  int tile_begin, tile_end;
  int group, group_rows;  // which wave-group this item belongs to and how many cluster rows it has
};

// Work items are enumerated per cluster: item t names a cluster row (`cluster`
// consecutive 128-query blocks, one per CTA of the cluster) and a chunk.
__host__ __device__ __forceinline__ bool get_item(const TopkPlan &p, int t, int cta_rank, Item *it) {
  const int per_group = p.group_mc * p.chunks;
  const int g = t / per_group;
  const int rem = t - g * per_group;
  const int m0 = g * p.group_mc;
  const int gm = p.group_mc < p.mc_blocks - m0 ? p.group_mc : p.mc_blocks - m0;
  if (gm <= 0 || rem >= gm * p.chunks) return false;
  const int c = rem / gm;
  it->group = g;
  it->group_rows = gm;
  it->m = ((m0 + (rem - c * gm) + p.rot) % p.mc_blocks) * p.cluster + cta_rank;  // may be a phantom block past m_blocks
  it->chunk = c;
  it->tile_begin = static_cast<int>(static_cast<long long>(c) * p.n_tiles / p.chunks);
  it->tile_end = static_cast<int>(static_cast<long long>(c + 1) * p.n_tiles / p.chunks);
  return true;
}

// The i-th item of a cluster.  Static schedule: round-robin over the item list (whole waves of
// co-resident groups, see make_topk_plan).  Dynamic schedule (plan.dyn_slots): the leader CTA's
// producer thread takes the next item from an atomic counter and publishes it in the cluster's
// slot i; every other role of both CTAs (MMA issuer, epilogue warps, the second CTA's producer)
// reads that slot when it gets there -- the producer runs a pipeline depth ahead, so the slot is
// normally filled long before.  Returns -1 when the cluster is done.
__device__ __forceinline__ int next_item(const TopkPlan &plan, int cluster_id, int num_clusters, int i, int total_items,
                                         bool grab) {
  if (plan.dyn_slots == nullptr) {
    const long long t = cluster_id + static_cast<long long>(i) * num_clusters;
    return t < total_items ? static_cast<int>(t) : -1;
  }
  unsigned *slot = plan.dyn_slots + static_cast<size_t>(cluster_id) * plan.dyn_cap + i;
  unsigned v;
  if (grab) {
    const unsigned t = atomicAdd(plan.dyn_counter, 1u);
    v = t < static_cast<unsigned>(total_items) ? t + 2u : 1u;
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(slot), "r"(v) : "memory");
  } else {
    const long long t0 = clock64();
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(slot) : "memory");
      if (v != 0) break;
      if (clock64() - t0 > 8000000000LL) __trap();   // a lost producer: fail, never hang
      __nanosleep(100);
    }
  }
  return v == 1u ? -1 : static_cast<int>(v - 2u);
}

__device__ __forceinline__ unsigned long long shfl_xor_u64(unsigned long long v, int mask) {
  return __shfl_xor_sync(0xFFFFFFFFu, v, mask);
}

// Bitonic sort of 64 keys held two per lane: `a` is element `lane`, `b` is
// element `lane + 32`.  Ascending; afterwards `a` holds the 32 smallest in lane
// order.
__device__ __forceinline__ void warp_sort64(unsigned long long &a, unsigned long long &b, int lane) {
#pragma unroll
  for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j == 32) {
        unsigned long long lo = a < b ? a : b, hi = a < b ? b : a;
        a = lo;
        b = hi;
      } else {
        unsigned long long pa = shfl_xor_u64(a, j), pb = shfl_xor_u64(b, j);
        const bool lower = (lane & j) == 0;
        const bool up_a = (lane & k) == 0;          // element index = lane
        const bool up_b = ((lane + 32) & k) == 0;   // element index = lane + 32
        unsigned long long mn = a < pa ? a : pa, mx = a < pa ? pa : a;
        a = (lower == up_a) ? mn : mx;
        mn = b < pb ? b : pb;
        mx = b < pb ? pb : b;
        b = (lower == up_b) ? mn : mx;
      }
    }
  }
}

// Warp-cooperative compaction of one row's list: sort, keep the KP smallest,
// publish the new count and threshold to the lane that owns the row.  With
// `emit` the sorted keys are also written to global memory.  The sort is kept
// out of line: it is the bulk of the epilogue's code and is called from several
// places; one copy keeps the epilogue within the instruction cache.
__device__ __noinline__ unsigned long long compact_sorted(uint32_t row_addr, int rot, int rcnt, int lane,
                                                          unsigned long long *emit) {
  unsigned long long a = kKeyInvalid, b = kKeyInvalid;
  const uint32_t addr_a = row_addr + (((lane + rot) & (CAP - 1)) << 3);
  const uint32_t addr_b = row_addr + (((lane + 32 + rot) & (CAP - 1)) << 3);
  if (lane < rcnt) asm volatile("ld.shared.b64 %0, [%1];" : "=l"(a) : "r"(addr_a) : "memory");
  if (lane + 32 < rcnt) asm volatile("ld.shared.b64 %0, [%1];" : "=l"(b) : "r"(addr_b) : "memory");
  warp_sort64(a, b, lane);
  asm volatile("st.shared.b64 [%0], %1;" ::"r"(addr_a), "l"(a) : "memory");
  if (emit != nullptr) emit[lane] = a;
  __syncwarp();
  return __shfl_sync(0xFFFFFFFFu, a, KP - 1);
}
__device__ __forceinline__ void compact_row(uint32_t row_addr, int rot, int r, int lane, int &cnt, float &tau,
                                            unsigned long long *emit) {
  const int rcnt = __shfl_sync(0xFFFFFFFFu, cnt, r);
  const unsigned long long last = compact_sorted(row_addr, rot, rcnt, lane, emit);
  if (lane == r) {
    cnt = rcnt < KP ? rcnt : KP;
    tau = (rcnt >= KP) ? key_score(last) : std::numeric_limits<float>::infinity();
  }
}

// Smem layout.  `STG` pipeline stages of (query tile, data tile[ half]) precede the
// per-row candidate lists, the staged data norms and the barriers.
template <int STG, int B_STAGE_BYTES, int LISTS = 1>
struct Layout {
  static constexpr int kOffA = 0;
  static constexpr int kOffB = kOffA + STG * A_BYTES;
  static constexpr int kOffList = kOffB + STG * B_STAGE_BYTES;
  static constexpr int kOffXn = kOffList + LISTS * BM * CAP * 8;
  static constexpr int kOffBar = kOffXn + 2 * BN * 4;
  static constexpr int kBytes = kOffBar + 256 + 1024;  // + alignment slack
};
using Layout1 = Layout<STAGES, B_BYTES>;            // one CTA per 128 x 256 tile
constexpr int STAGES2 = 4;
using Layout2 = Layout<STAGES2, B_BYTES / 2>;       // CTA pair per 256 x 256 tile
// short rows: the pair kernel with two sets of epilogue warps and row lists; a stage is a whole
// tile there (one or two k-blocks) and the epilogue is what the tile waits for, so two stages do
using Layout3 = Layout<2, B_BYTES / 2, 2>;

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

// Epilogue shared by both kernels: 4 warps, one query row per thread, 256
// accumulator columns per tile.  `tempty_addr` is the barrier (cluster address)
// on which each warp signals that the accumulator stage may be overwritten.
template <bool REMOTE_EMPTY>
__device__ __forceinline__ void epilogue_loop(const TopkPlan &plan, uint8_t *smem, int off_list, int off_xn,
                                              uint32_t tmem_base, uint64_t *tfull_bar, uint64_t *tempty_bar,
                                              int cluster_id, int num_clusters, int cta_rank, int total_items,
                                              const float *__restrict__ xnorm2,
                                              unsigned long long *__restrict__ cand, int half = 0, int nhalves = 1) {
  // `half` of `nhalves`: with two sets of epilogue warps (short rows) set h scans columns
  // [h * 128, h * 128 + 128) of every tile into its own row lists (`off_list` is the set's) and
  // emits list 2 * chunk + h; the staging of the data norms is shared by all epilogue threads
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quarter = warp & 3;              // tensor-memory lane quarter this warp may read
  const int row_in_tile = quarter * 32 + lane;
  const uint32_t lists_addr = smem_u32(smem + off_list);  // [BM][CAP] keys, shared-space address
  const uint32_t keys_base = lists_addr + row_in_tile * CAP * 8;  // this row's list, shared-space address
  float *xn_all = reinterpret_cast<float *>(smem + off_xn);
  const int e_tid = half == 0 ? threadIdx.x - 64 : threadIdx.x - 96;   // (warp 6 is the communication warp)
  const int epi_threads = EPI_THREADS * nhalves;
  const int col_begin = half * (BN / nhalves), col_end = col_begin + BN / nhalves;
  const float inf = std::numeric_limits<float>::infinity();
  int cnt = 0;
  float tau = inf;
  uint32_t acc_it = 0;
  uint32_t uses[2] = {0, 0};  // how often each accumulator stage has been drained
  long long prof_wait = 0, prof_stage = 0, prof_ld = 0, prof_score = 0, prof_compact = 0, prof_flush = 0;
  const bool prof = plan.prof != nullptr;
  const long long prof_t0 = prof ? clock64() : 0;

  for (int i = 0;; ++i) {
    int t = lane == 0 ? next_item(plan, cluster_id, num_clusters, i, total_items, false) : 0;
    t = __shfl_sync(0xFFFFFFFFu, t, 0);
    if (t < 0) break;
    Item it;
    if (!get_item(plan, t, cta_rank, &it)) continue;
    for (int tile = it.tile_begin; tile < it.tile_end; ++tile, ++acc_it) {
      // stages alternate tile by tile; the wide pair kernel restarts at stage 0 with every item
      // (its odd tiles own stage 1, and an item with an odd tile count leaves that one unused)
      const uint32_t acc = plan.wide ? static_cast<uint32_t>((tile - it.tile_begin) & 1) : (acc_it & 1);
      const uint32_t acc_phase = (acc ? uses[1] : uses[0]) & 1;
      if (acc) ++uses[1]; else ++uses[0];
      float *xn = xn_all + (acc_it & 1) * BN;  // (not `acc`: two tiles in a row can share a stage across items)
      // The data norms of the tile are staged while its MMAs still run.  When the rows (and their
      // norms) are being gathered from peer GPUs by this very kernel, one thread first makes sure the
      // tile has arrived (the counter the TMA producer waited on before it loaded the tile, so this
      // hardly ever spins); staging without that check raced with the gather.
      const long long ps0 = prof ? clock64() : 0;
      if (plan.gate_cnt != nullptr) {
        if (e_tid == 0) {
          const int rows_in_tile = min(BN, plan.n - tile * BN);
          const unsigned target = plan.gate_epoch * static_cast<unsigned>(rows_in_tile);
          const long long t0 = clock64();
          while (static_cast<int>(ld_acquire_gpu_u32(plan.gate_cnt + tile) - target) < 0)
            if (clock64() - t0 > 4000000000LL) break;   // the producer's own watchdog reports a lost peer
        }
        named_barrier_sync(1, epi_threads);
      }
      for (int i = e_tid; i < BN; i += epi_threads) {
        const int col = tile * BN + i;
        xn[i] = col < plan.n ? __ldcg(xnorm2 + col) : inf;
      }
      named_barrier_sync(1, epi_threads);
      const long long pw0 = prof ? clock64() : 0;
      if (prof) prof_stage += pw0 - ps0;
      mbar_wait(&tfull_bar[acc], acc_phase);
      if (prof) prof_wait += clock64() - pw0;
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      // (one tcgen05.ld in flight at a time: overlapping the load of the next 32 columns with
      // the scoring of the current ones measured 25-80 % slower, for the MMA as well)
      const uint32_t xn_addr = smem_u32(xn);
#pragma unroll 1
      for (int c0 = col_begin; c0 < col_end; c0 += 32) {
        uint32_t v[32];
        float x[32];
        const long long pl0 = prof ? clock64() : 0;
        tmem_ld_32x32(t_row + c0, v);
#pragma unroll
        for (int i = 0; i < 8; ++i)  // the norms of these 32 rows (same address in every lane: a broadcast)
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(x[4 * i]), "=f"(x[4 * i + 1]), "=f"(x[4 * i + 2]), "=f"(x[4 * i + 3])
                       : "r"(xn_addr + (c0 + 4 * i) * 4));
        tmem_ld_wait();
        if (prof) {  // the load is only complete once a register it fills can be read
          uint32_t sink;
          asm volatile("mov.b32 %0, %1;" : "=r"(sink) : "r"(v[31]));
          asm volatile("" ::"r"(sink));
        }
        const long long pl1 = prof ? clock64() : 0;
        if (prof) prof_ld += pl1 - pl0;
        long long pc_blk = 0;
        // all 32 scores at once (independent FMAs), then one octet at a time: the octet's minimum
        // against the threshold decides, for the whole warp, whether its append code runs at all
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = fmaf(-2.0f, __uint_as_float(v[j]), x[j]);
        const int id0 = plan.row_base + tile * BN + c0;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
          const float *sc = x + 8 * o;
          const float mn = fminf(fminf(fminf(sc[0], sc[1]), fminf(sc[2], sc[3])),
                                 fminf(fminf(sc[4], sc[5]), fminf(sc[6], sc[7])));
          if (!__any_sync(0xFFFFFFFFu, mn < tau)) continue;
          // Room for the 8 appends an octet can make.  Compacting only when a list is nearly
          // full (not whenever it holds more than KP keys) costs a slightly stale threshold
          // and saves most of the sorts: a row appends ~250 keys per 30 000 columns, mostly one
          // at a time, and a sort per append was the bulk of the epilogue's instructions.
          unsigned need = __ballot_sync(0xFFFFFFFFu, cnt > CAP - 8);
          if (need) {
            const long long pc0 = prof ? clock64() : 0;
            while (need) {
              const int r = __ffs(need) - 1;
              need &= need - 1;
              compact_row(lists_addr + (quarter * 32 + r) * CAP * 8, quarter * 32 + r, r, lane, cnt, tau, nullptr);
            }
            if (prof) pc_blk += clock64() - pc0;
          }
          // Branch-free appends: every lane runs the same 8 predicated stores.  The slot of
          // store j is the list's fill plus the hits before j -- a small adder tree, so that the
          // eight address computations and stores are independent of each other (a running
          // `cnt += hit` chained them: ~200 cycles of latency per octet, for the whole warp, and
          // with short rows about every second octet has a hit in one of its 32 rows).
          // Hits count in units of 8 (bytes per key); a row's list is 512-byte aligned.
          int h[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) h[j] = sc[j] < tau ? 8 : 0;
          const int p2 = h[0] + h[1], p4 = p2 + (h[2] + h[3]), p6 = p4 + (h[4] + h[5]);
          const int pre[8] = {0, h[0], p2, p2 + h[2], p4, p4 + h[4], p6, p6 + h[6]};
          const int s8 = (cnt + row_in_tile) << 3;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t addr = keys_base | (static_cast<uint32_t>(s8 + pre[j]) & ((CAP - 1) << 3));
            const uint32_t b = __float_as_uint(sc[j]);
            const uint32_t ob = b ^ (static_cast<uint32_t>(static_cast<int>(b) >> 31) | 0x80000000u);  // float_to_ordered
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "setp.lt.f32 p, %0, %1;\n\t"
                "@p st.shared.v2.b32 [%2], {%3, %4};\n\t}" ::"f"(sc[j]),
                "f"(tau), "r"(addr), "r"(id0 + 8 * o + j), "r"(ob)
                : "memory");
          }
          cnt += (p6 + (h[6] + h[7])) >> 3;
        }
        if (prof) {
          prof_compact += pc_blk;
          prof_score += clock64() - pl1 - pc_blk;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (REMOTE_EMPTY) mbar_arrive_remote(smem_u32(&tempty_bar[acc]) & 0xFEFFFFFFu);  // leader CTA's barrier
        else mbar_arrive(&tempty_bar[acc]);
      }
    }
    // flush this (row block, chunk): sorted kp keys per row
    __syncwarp();
    const long long pf0 = prof ? clock64() : 0;
    for (int r = 0; r < 32; ++r) {
      const long long qrow = static_cast<long long>(it.m) * BM + quarter * 32 + r;
      unsigned long long *out =
          qrow < plan.q_rows ? cand + (qrow * plan.cand_chunks + plan.chunk_base + it.chunk * nhalves + half) * KP
                             : nullptr;
      compact_row(lists_addr + (quarter * 32 + r) * CAP * 8, quarter * 32 + r, r, lane, cnt, tau, out);
    }
    cnt = 0;
    tau = inf;
    if (prof) prof_flush += clock64() - pf0;
  }
  if (prof && quarter == 0 && lane == 0) {
    unsigned long long *o = plan.prof + blockIdx.x * 16;
    o[6] = prof_wait;
    o[7] = clock64() - prof_t0;
    o[8] = prof_stage;
    o[9] = prof_ld;
    o[10] = prof_score;
    o[11] = prof_compact;
    o[12] = prof_flush;
  }
}

// ---------------------------------------------------------------------------
// Variant 1: one CTA per 128 x 256 tile (tcgen05 cta_group::1); optional cluster
// whose CTAs share each data tile through TMA multicast.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS, 1)
dist_topk_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_x,
                 const float *__restrict__ xnorm2, const TopkPlan plan, unsigned long long *__restrict__ cand) {
  using L = Layout1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L::kOffBar);
  uint64_t *full_bar = bars;                  // [STAGES]
  uint64_t *empty_bar = bars + STAGES;        // [STAGES]
  uint64_t *tfull_bar = bars + 2 * STAGES;    // [2]
  uint64_t *tempty_bar = bars + 2 * STAGES + 2;  // [2]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int csize = plan.cluster;
  const int cta_rank = csize > 1 ? static_cast<int>(cluster_ctarank()) : 0;
  const int cluster_id = blockIdx.x / csize;
  const int num_clusters = gridDim.x / csize;
  const uint16_t cmask = static_cast<uint16_t>((1u << csize) - 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_x);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], csize);  // every CTA of the cluster must release the stage
      }
      for (int s = 0; s < 2; ++s) {
        mbar_init(&tfull_bar[s], 1);
        mbar_init(&tempty_bar[s], 4);  // one arrive per epilogue warp
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  if (csize > 1) cluster_sync_all();  // peers' barriers must exist before multicast traffic starts
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int per_group = plan.group_mc * plan.chunks;
  const int total_items = ((plan.mc_blocks + plan.group_mc - 1) / plan.group_mc) * per_group;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint64_t pol_q = plan.hint_q == 2 ? kL2EvictLast : (plan.hint_q == 1 ? kL2EvictFirst : kL2EvictNormal);
      const uint64_t pol_x = plan.hint_x == 2 ? kL2EvictLast : (plan.hint_x == 1 ? kL2EvictFirst : kL2EvictNormal);
      const int slice_rows = BN / csize, slice_bytes = slice_rows * BK * 2;
      for (int t = cluster_id; t < total_items; t += num_clusters) {
        Item it;
        if (!get_item(plan, t, cta_rank, &it)) continue;
        for (int tile = it.tile_begin; tile < it.tile_end; ++tile) {
          for (int kb = 0; kb < plan.k_blocks; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
            // query tiles are re-read for every data tile of the sweep: keep them in L2
            tma_load_2d_hint(smem + L::kOffA + stage * A_BYTES, &tm_q, &full_bar[stage], kb * BK, it.m * BM, pol_q);
            if (csize == 1) {
              tma_load_2d_hint(smem + L::kOffB + stage * B_BYTES, &tm_x, &full_bar[stage], kb * BK, tile * BN, pol_x);
            } else {
              // this CTA fetches one slice of the data tile and multicasts it to the cluster
              tma_load_2d_mcast(smem + L::kOffB + stage * B_BYTES + cta_rank * slice_bytes, &tm_x, &full_bar[stage],
                                kb * BK, tile * BN + cta_rank * slice_rows, cmask, pol_x);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------- MMA issuer -------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16(BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_it = 0;
      for (int t = cluster_id; t < total_items; t += num_clusters) {
        Item it;
        if (!get_item(plan, t, cta_rank, &it)) continue;
        for (int tile = it.tile_begin; tile < it.tile_end; ++tile, ++acc_it) {
          const uint32_t acc = acc_it & 1, acc_phase = (acc_it >> 1) & 1;
          mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * BN;
          for (int kb = 0; kb < plan.k_blocks; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint64_t da = make_smem_desc_sw128(smem_u32(smem + L::kOffA + stage * A_BYTES));
            const uint64_t db = make_smem_desc_sw128(smem_u32(smem + L::kOffB + stage * B_BYTES));
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk)
              umma_f16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
            // frees the smem stage (in every CTA of the cluster) once these MMAs retire
            if (csize == 1) umma_commit(&empty_bar[stage]);
            else umma_commit_mcast(&empty_bar[stage], cmask);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          umma_commit(&tfull_bar[acc]);      // accumulator complete -> epilogue
        }
      }
    }
  } else {
    epilogue_loop<false>(plan, smem, L::kOffList, L::kOffXn, tmem_base, tfull_bar, tempty_bar, cluster_id,
                         num_clusters, cta_rank, total_items, xnorm2, cand);
  }

  tc_fence_before();
  __syncthreads();
  if (csize > 1) cluster_sync_all();  // no CTA may leave while peers can still signal it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------
// Variant 2: CTA pair per 256 x 256 tile (tcgen05 cta_group::2).  Each CTA of the
// pair stages its own 128 query rows and one half (128 rows) of the data tile;
// the leader CTA issues the MMAs for both, so each SM's shared memory moves
// half as many operand bytes per flop -- what the 128 B/clk smem port needs for
// the tensor pipe to run at full rate.  Accumulator rows live in each CTA's own
// tensor memory, so the epilogue is the same as in variant 1.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d_pair(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0,
                                                 int c1, uint64_t policy) {
  // completion bytes are credited to the barrier at this offset in the leader CTA
  const uint32_t leader_bar = smem_u32(bar) & 0xFEFFFFFFu;
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1),
      "l"(policy)
      : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t *bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
      : "memory");
}

constexpr int THREADS2 = THREADS + 32;  // + one communication warp
constexpr int THREADS3 = THREADS2 + EPI_THREADS;  // + a second set of epilogue warps (short rows, MODE 3)

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Communication warp (K4): all-gather of the packed query rows, fused into the
// GEMM.  Rows are taken in the order the tiles will need them; each finished row
// bumps its block's counter, which the TMA producers poll before touching the
// block.  Afterwards the fp64 rows, norms and k of the remote slices are pulled
// for the re-rank kernel (no flags needed: it runs after this kernel).
__device__ void gather_warp(const TopkPlan &plan, const GatherArgs &ga, int lane) {
  const int G = gridDim.x, b = blockIdx.x;
  unsigned ready_mask = 1u << ga.rank;
  bool dead = false;
  auto wait_peer = [&](int owner) {
    if ((ready_mask >> owner) & 1u) return;
    const unsigned *flag = reinterpret_cast<const unsigned *>(ga.peer_base[ga.rank] + ga.off_pack_flags) + owner;
    const long long t0 = clock64();
    while (static_cast<int>(ld_acquire_sys(flag) - ga.epoch) < 0) {
      if (clock64() - t0 > ga.timeout_cycles) {  // watchdog: report and stop waiting
        if (lane == 0 && ga.status) atomicExch(ga.status, 100 + owner);
        dead = true;
        break;
      }
      __nanosleep(100);
    }
    ready_mask |= 1u << owner;
  };
  const int vec_per_row = plan.a_pad / 8;  // int4 = 8 halves
  // rows travel in groups of RB consecutive rows (one owner: slices are whole 256-row tiles), so
  // that 32 lanes x 8 independent 16-byte loads = 4 KB are in flight per round trip for short rows too
  int RB = 1;
  while (RB < 8 && 2 * RB * vec_per_row <= 256) RB <<= 1;
  const int gpb = BM / RB;  // groups per 128-row block
  auto copy_vec = [&](const int4 *src, int4 *dst, int count) {  // contiguous, 16-byte units
    for (int v0 = 0; v0 < count; v0 += 32 * 8) {
      int4 tmp[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < count) tmp[u] = src[v];
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < count) dst[v] = tmp[u];
      }
    }
  };
  unsigned char *own = ga.peer_base[ga.rank];
  const long long total_groups = static_cast<long long>(plan.mc_blocks) * plan.cluster * gpb;
  for (long long seq = b; seq < total_groups; seq += G) {
    const int o = static_cast<int>(seq / gpb), g = static_cast<int>(seq - static_cast<long long>(o) * gpb);
    const int cr = o / plan.cluster, j = o - cr * plan.cluster;
    const int m = ((cr + plan.rot) % plan.mc_blocks) * plan.cluster + j;
    const long long grow = static_cast<long long>(m) * BM + g * RB;
    const int owner = static_cast<int>(grow / ga.slice_rows);
    if (owner != ga.rank && owner < ga.world) {
      wait_peer(owner);
      if (!dead)
        copy_vec(reinterpret_cast<const int4 *>(ga.peer_base[owner] + ga.off_qh) + grow * vec_per_row,
                 reinterpret_cast<int4 *>(own + ga.off_qh) + grow * vec_per_row, RB * vec_per_row);
    }
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      atomicAdd(ga.rb_count + m, static_cast<unsigned>(RB));
    }
  }
  // fp64 rows + per-query scalars of the remote slices, for the re-rank kernel: every slice is one
  // contiguous span (slice_rows is a multiple of 256), copied in 4 KB pieces by all communication warps
  for (int pj = 1; pj < ga.world && !dead; ++pj) {
    const int owner = (ga.rank + pj) % ga.world;
    wait_peer(owner);
    if (dead) break;
    const unsigned char *peer = ga.peer_base[owner];
    const long long r0 = static_cast<long long>(owner) * ga.slice_rows;
    const long long nvec = static_cast<long long>(ga.slice_rows) * ga.a / 2;   // slice_rows is even
    const int4 *s4 = reinterpret_cast<const int4 *>(reinterpret_cast<const double *>(peer + ga.off_q64) + r0 * ga.a);
    int4 *d4 = reinterpret_cast<int4 *>(reinterpret_cast<double *>(own + ga.off_q64) + r0 * ga.a);
    for (long long v0 = static_cast<long long>(b) * 256; v0 < nvec; v0 += static_cast<long long>(G) * 256) {
      const long long left = nvec - v0;
      copy_vec(s4 + v0, d4 + v0, left < 256 ? static_cast<int>(left) : 256);
    }
    for (long long i = r0 + static_cast<long long>(b) * 32 + lane; i < r0 + ga.slice_rows; i += static_cast<long long>(G) * 32) {
      reinterpret_cast<float *>(own + ga.off_qn)[i] = reinterpret_cast<const float *>(peer + ga.off_qn)[i];
      reinterpret_cast<float *>(own + ga.off_qe)[i] = reinterpret_cast<const float *>(peer + ga.off_qe)[i];
      reinterpret_cast<int *>(own + ga.off_k)[i] = reinterpret_cast<const int *>(peer + ga.off_k)[i];
    }
  }
}

// Communication warp for the data operand (fused row all-gather).  Rows are taken tile by
// tile in sweep order, the chunks interleaved; a finished row bumps its tile's counter.
// Then the fp64 rows and labels of the other parts follow, starting with a different peer
// on every rank.
__device__ void data_gather_warp(const TopkPlan &plan, const DataGatherArgs &dg, int lane) {
  const int G = gridDim.x, b = blockIdx.x;
  unsigned ready_mask = 1u << dg.rank;
  bool dead = false;
  auto wait_peer = [&](int owner) {
    if ((ready_mask >> owner) & 1u) return;
    const unsigned *flag = reinterpret_cast<const unsigned *>(dg.peer_base[dg.rank] + dg.off_pack_flags) + owner;
    const long long t0 = clock64();
    while (static_cast<int>(ld_acquire_sys(flag) - dg.epoch) < 0) {
      if (clock64() - t0 > dg.timeout_cycles) {
        if (lane == 0 && dg.status) atomicExch(dg.status, 400 + owner);
        dead = true;
        break;
      }
      __nanosleep(100);
    }
    ready_mask |= 1u << owner;
  };
  auto owner_of = [&](int row) {
    int o = 0;
    while (o + 1 < dg.world && row >= dg.part_begin[o + 1]) ++o;
    return o;
  };
  const int vec_per_row = plan.a_pad / 8;  // int4 = 8 halves
  // 32 lanes x 8 independent 16-byte loads = 4 KB in flight per warp and round trip: rows are
  // taken in groups of RB consecutive rows so that short rows fill that window too (one row of
  // a = 256 is only 512 bytes; a round trip over NVLink costs microseconds whatever its size)
  int RB = 1;
  while (RB < 8 && 2 * RB * vec_per_row <= 256) RB <<= 1;
  const int gpt = BN / RB;  // groups per tile
  auto copy_vec = [&](const int4 *src, int4 *dst, int count) {  // contiguous, 16-byte units
    for (int v0 = 0; v0 < count; v0 += 32 * 8) {
      int4 tmp[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < count) tmp[u] = src[v];
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < count) dst[v] = tmp[u];
      }
    }
  };
  const int per_chunk = (plan.n_tiles + plan.chunks - 1) / plan.chunks;
  const long long total = static_cast<long long>(per_chunk) * plan.chunks * gpt;
  unsigned char *own = dg.peer_base[dg.rank];
  for (long long seq = b; seq < total; seq += G) {
    const int g = static_cast<int>(seq % gpt);
    const long long ts = seq / gpt;
    const int c = static_cast<int>(ts % plan.chunks), j = static_cast<int>(ts / plan.chunks);
    const int tb = static_cast<int>(static_cast<long long>(c) * plan.n_tiles / plan.chunks);
    const int te = static_cast<int>(static_cast<long long>(c + 1) * plan.n_tiles / plan.chunks);
    const int tile = tb + j;
    if (tile >= te) continue;
    const int row0 = tile * BN + g * RB;
    const int rows = min(RB, plan.n - row0);
    if (rows <= 0) continue;
    // a group never straddles more than two parts; copy it part by part (usually in one piece)
    for (int r = row0; r < row0 + rows;) {
      const int owner = owner_of(r);
      const int r_end = min(row0 + rows, dg.part_begin[owner + 1]);
      if (owner != dg.rank) {
        wait_peer(owner);
        if (!dead) {
          const unsigned char *peer = dg.peer_base[owner];
          copy_vec(reinterpret_cast<const int4 *>(peer + dg.off_xh) + static_cast<long long>(r) * vec_per_row,
                   reinterpret_cast<int4 *>(own + dg.off_xh) + static_cast<long long>(r) * vec_per_row,
                   (r_end - r) * vec_per_row);
          if (lane < r_end - r) {
            reinterpret_cast<float *>(own + dg.off_xn)[r + lane] = reinterpret_cast<const float *>(peer + dg.off_xn)[r + lane];
            reinterpret_cast<float *>(own + dg.off_xe)[r + lane] = reinterpret_cast<const float *>(peer + dg.off_xe)[r + lane];
          }
        }
      }
      r = r_end;
    }
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      atomicAdd(dg.tile_count + tile, static_cast<unsigned>(rows));
    }
  }
  // fp64 rows + labels of the other parts, for the re-rank kernel that follows: each part is one
  // contiguous span, copied in 4 KB pieces spread over all communication warps
  for (int pj = 1; pj < dg.world && !dead; ++pj) {
    const int owner = (dg.rank + pj) % dg.world;
    wait_peer(owner);
    if (dead) break;
    const unsigned char *peer = dg.peer_base[owner];
    const int units = dg.x_units > 0 ? dg.x_units : dg.a;                            // 8-byte units per row (fp64 or decimal6 rows)
    const long long e0 = static_cast<long long>(dg.part_begin[owner]) * units;      // first unit of the part
    const long long e1 = static_cast<long long>(dg.part_begin[owner + 1]) * units;  // one past its last
    const double *src = reinterpret_cast<const double *>(peer + dg.off_x64);
    double *dst = reinterpret_cast<double *>(own + dg.off_x64);
    long long lo = e0, hi = e1;
    if (lo & 1) { if (b == 0 && lane == 0 && lo < hi) dst[lo] = src[lo]; ++lo; }      // 16-byte alignment of the bulk
    if ((hi & 1) && hi > lo) { --hi; if (b == 0 && lane == 1) dst[hi] = src[hi]; }
    const long long nvec = hi > lo ? (hi - lo) / 2 : 0;
    const int4 *s4 = reinterpret_cast<const int4 *>(src + lo);
    int4 *d4 = reinterpret_cast<int4 *>(dst + lo);
    for (long long v0 = static_cast<long long>(b) * 256; v0 < nvec; v0 += static_cast<long long>(G) * 256) {
      const long long left = nvec - v0;
      copy_vec(s4 + v0, d4 + v0, left < 256 ? static_cast<int>(left) : 256);
    }
    const int *sl = reinterpret_cast<const int *>(peer + dg.off_labels);
    int *dl = reinterpret_cast<int *>(own + dg.off_labels);
    for (long long i = dg.part_begin[owner] + static_cast<long long>(b) * 32 + lane; i < dg.part_begin[owner + 1];
         i += static_cast<long long>(G) * 32)
      dl[i] = sl[i];
  }
}

// Skewed wide schedule (MODE 2): which k-step an item is at decides what is in flight.  Tile j
// of the item runs on accumulator j & 1 during steps [j * H, j * H + K); H = (K + gap) / 2, so
// the two accumulators are half a period apart, each has `gap` steps between the end of one
// tile and the start of its next one (its drain), and at most two tiles are active at a step.
__device__ __forceinline__ int skew_first_tile(int s, int K, int H) {
  const int j = (s - K + H) / H;  // ceil((s - K + 1) / H) for s >= K - H, else <= 0
  return s - K + 1 <= 0 ? 0 : j;
}

template <int MODE>
__global__ void __launch_bounds__(MODE == 3 ? THREADS3 : THREADS2, 1)
dist_topk_pair_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_x,
                      const float *__restrict__ xnorm2, const TopkPlan plan,
                      unsigned long long *__restrict__ cand, const GatherArgs ga, const DataGatherArgs dg) {
  // WIDE: every stage holds one query tile and this CTA's halves of TWO data tiles, and both
  // accumulator stages are filled side by side (256 x 512 per pair and k-loop).  The query
  // operand then crosses L2 -> SM once per two data tiles: 48 instead of 64 bytes per SM and
  // clock at full tensor rate, against an L2 that delivers ~43 (measured 8.7 TB/s at 72 %).
  // The price is that the two stages drain while the MMA waits.
  // SKEW (MODE 2) keeps the wide tile's operand sharing but never lets the tensor pipe wait for a
  // drain: the two accumulators run half a period apart, each tile walks the k-blocks
  // cyclically from wherever the (continuous) query stream is when it starts, and a tile's
  // accumulator rests for `gap` steps -- while the other one keeps the MMA busy -- before its
  // next tile begins.  The query tile of a step is still shared by both active data tiles.
  // HALVES (MODE 3, short rows): the narrow tile with a second set of four epilogue warps.  With one
  // or two k-blocks per tile the kernel is bound by the epilogue (the MMA waits for a free
  // accumulator 90 % of the time at A = 64, profiles/topk_prof_short_r2.log), and a warp can only
  // read its own TMEM lane quarter -- so the second set shares the rows and takes the other 128
  // columns of every tile, with row lists of its own (two lists per query and chunk).
  constexpr bool WIDE = MODE == 1 || MODE == 2;
  constexpr bool SKEW = MODE == 2;
  constexpr bool HALVES = MODE == 3;
  constexpr int NST = WIDE ? STAGES : (HALVES ? 2 : STAGES2);
  constexpr int BH_BYTES = B_BYTES / 2;  // this CTA's half of one data tile
  constexpr int BS_BYTES = WIDE ? B_BYTES : BH_BYTES;  // data bytes per stage
  using L = Layout<NST, BS_BYTES, HALVES ? 2 : 1>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L::kOffBar);
  uint64_t *full_bar = bars;                    // [NST]  (the leader's copy is the live one)
  uint64_t *empty_bar = bars + NST;         // [NST]  per CTA, released by the leader's commits
  uint64_t *tfull_bar = bars + 2 * NST;     // [2]        per CTA
  uint64_t *tempty_bar = bars + 2 * NST + 2;  // [2]      leader's copy collects both epilogues
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * NST + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int cta_rank = static_cast<int>(cluster_ctarank());  // 0 = leader
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_x);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < NST; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int s = 0; s < 2; ++s) {
        mbar_init(&tfull_bar[s], 1);
        mbar_init(&tempty_bar[s], HALVES ? 16 : 8);  // 4 (or 8) epilogue warps in each CTA of the pair
      }
      fence_barrier_init();
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int per_group = plan.group_mc * plan.chunks;
  const int total_items = ((plan.mc_blocks + plan.group_mc - 1) / plan.group_mc) * per_group;

  if (warp == 0) {
    // ---------------- TMA producer (both CTAs load their own operands) ----------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      long long pw_empty = 0, pw_sync = 0;
      const long long p_t0 = plan.prof ? clock64() : 0;
      const uint64_t pol_q = plan.hint_q == 2 ? kL2EvictLast : (plan.hint_q == 1 ? kL2EvictFirst : kL2EvictNormal);
      const uint64_t pol_x = plan.hint_x == 2 ? kL2EvictLast : (plan.hint_x == 1 ? kL2EvictFirst : kL2EvictNormal);
      for (int i = 0;; ++i) {
        const int t = next_item(plan, cluster_id, num_clusters, i, total_items, /*grab=*/cta_rank == 0);
        if (t < 0) break;
        Item it;
        if (!get_item(plan, t, cta_rank, &it)) continue;
        unsigned sync_epoch = 0;
        if (ga.world > 1) {
          // the block's packed rows may still be on their way from their owner
          const unsigned target = ga.epoch * BM;
          const long long t0 = clock64();
          while (static_cast<int>(ld_acquire_gpu(ga.rb_count + it.m) - target) < 0) {
            if (clock64() - t0 > ga.timeout_cycles) {
              if (ga.status) atomicExch(ga.status, 200);
              break;
            }
            __nanosleep(64);
          }
          asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy writes -> TMA (async proxy) reads
        }
        if constexpr (SKEW) {
          const int K = plan.k_blocks, H = (K + plan.gap) >> 1, T = it.tile_end - it.tile_begin;
          const int steps = (T - 1) * H + K;
          for (int st = 0; st < steps; ++st) {
            const int j0 = skew_first_tile(st, K, H);
            int j1 = st / H;
            if (j1 > T - 1) j1 = T - 1;
            const int kb = st % K;
            if (dg.world > 1 && j1 * H == st) {
              // a tile starts: its packed rows (and norms) may still be on their way
              const int tile = it.tile_begin + j1;
              const int rows_in_tile = min(BN, plan.n - tile * BN);
              const unsigned target = dg.epoch * static_cast<unsigned>(rows_in_tile);
              const long long t0 = clock64();
              while (static_cast<int>(ld_acquire_gpu(dg.tile_count + tile) - target) < 0) {
                if (clock64() - t0 > dg.timeout_cycles) {
                  if (dg.status) atomicExch(dg.status, 500);
                  break;
                }
                __nanosleep(64);
              }
              asm volatile("fence.proxy.async;" ::: "memory");
            }
            if (plan.sync && cta_rank == 0 && it.group_rows > 1 && (st % (K * plan.sync)) == 0) {
              // re-align the pairs that sweep this chunk (see the tile-ordered loop below)
              unsigned *cnt = plan.sync_cnt + (it.group * plan.chunks + it.chunk);
              atomicAdd(cnt, 1u);
              ++sync_epoch;
              const unsigned target = static_cast<unsigned>(it.group_rows) * sync_epoch;
              const long long t0 = clock64();
              while (static_cast<int>(ld_acquire_gpu(cnt) - target) < 0) {
                if (clock64() - t0 > 2000000000LL) {
                  atomicAdd(&g_sync_timeouts, 1u);
                  break;
                }
              }
              if (plan.prof) pw_sync += clock64() - t0;
            }
            const long long pe0 = plan.prof ? clock64() : 0;
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (plan.prof) pw_empty += clock64() - pe0;
            if (cta_rank == 0) mbar_expect_tx(&full_bar[stage], 2 * (A_BYTES + (j1 - j0 + 1) * BH_BYTES));
            tma_load_2d_pair(smem + L::kOffA + stage * A_BYTES, &tm_q, &full_bar[stage], kb * BK, it.m * BM, pol_q);
            for (int j = j0; j <= j1; ++j)   // slot = accumulator = parity of the tile within the item
              tma_load_2d_pair(smem + L::kOffB + stage * BS_BYTES + (j & 1) * BH_BYTES, &tm_x, &full_bar[stage], kb * BK,
                               (it.tile_begin + j) * BN + cta_rank * (BN / 2), pol_x);
            if (++stage == NST) { stage = 0; phase ^= 1; }
          }
          continue;
        }
        constexpr int TSTEP = WIDE ? 2 : 1;
        for (int tile = it.tile_begin; tile < it.tile_end; tile += TSTEP) {
          const bool two = WIDE && tile + 1 < it.tile_end;
          if (dg.world > 1) {
            // the tile's packed rows (and norms) may still be on their way from the rank that packed them
            for (int h = 0; h <= (two ? 1 : 0); ++h) {
              const int rows_in_tile = min(BN, plan.n - (tile + h) * BN);
              const unsigned target = dg.epoch * static_cast<unsigned>(rows_in_tile);
              const long long t0 = clock64();
              while (static_cast<int>(ld_acquire_gpu(dg.tile_count + tile + h) - target) < 0) {
                if (clock64() - t0 > dg.timeout_cycles) {
                  if (dg.status) atomicExch(dg.status, 500);
                  break;
                }
                __nanosleep(64);
              }
            }
            asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy writes -> TMA (async proxy) reads
          }
          for (int kb = 0; kb < plan.k_blocks; ++kb) {
            if (plan.sync && cta_rank == 0 && it.group_rows > 1 && kb == 0 &&
                (((tile - it.tile_begin) / TSTEP) % plan.sync) == 0) {
              // Re-align the pairs that sweep this chunk.  They read the same data tiles; L2
              // only holds a few k-blocks of every stream, so once the pairs drift apart each
              // one re-fetches the tile from HBM.  One counter per (group, chunk): arrive,
              // then wait until all of the group's pairs have reached this point of the sweep.
              unsigned *cnt = plan.sync_cnt + (it.group * plan.chunks + it.chunk);
              atomicAdd(cnt, 1u);
              ++sync_epoch;
              const unsigned target = static_cast<unsigned>(it.group_rows) * sync_epoch;
              const long long t0 = clock64();
              while (static_cast<int>(ld_acquire_gpu(cnt) - target) < 0) {
                if (clock64() - t0 > 2000000000LL) {  // never hang on a lost peer, but say so
                  atomicAdd(&g_sync_timeouts, 1u);
                  break;
                }
              }
              if (plan.prof) pw_sync += clock64() - t0;
            }
            const long long pe0 = plan.prof ? clock64() : 0;
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (plan.prof) pw_empty += clock64() - pe0;
            if (cta_rank == 0) mbar_expect_tx(&full_bar[stage], 2 * (A_BYTES + (two ? 2 : 1) * BH_BYTES));
            tma_load_2d_pair(smem + L::kOffA + stage * A_BYTES, &tm_q, &full_bar[stage], kb * BK, it.m * BM, pol_q);
            tma_load_2d_pair(smem + L::kOffB + stage * BS_BYTES, &tm_x, &full_bar[stage], kb * BK,
                             tile * BN + cta_rank * (BN / 2), pol_x);
            if (two)
              tma_load_2d_pair(smem + L::kOffB + stage * BS_BYTES + BH_BYTES, &tm_x, &full_bar[stage], kb * BK,
                               (tile + 1) * BN + cta_rank * (BN / 2), pol_x);
            if (++stage == NST) { stage = 0; phase ^= 1; }
          }
        }
      }
      if (plan.prof) {
        plan.prof[blockIdx.x * 16 + 3] = clock64() - p_t0;
        plan.prof[blockIdx.x * 16 + 4] = pw_empty;
        plan.prof[blockIdx.x * 16 + 5] = pw_sync;
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (leader CTA only) ----------------
    if (lane == 0 && cta_rank == 0) {
      constexpr uint32_t idesc = make_idesc_f16(2 * BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_it = 0;
      uint32_t uses[2] = {0, 0};  // fills of each accumulator stage (same count as the epilogue keeps)
      long long mw_full = 0, mw_tempty = 0;
      const long long m_t0 = plan.prof ? clock64() : 0;
      for (int i = 0;; ++i) {
        const int t = next_item(plan, cluster_id, num_clusters, i, total_items, false);
        if (t < 0) break;
        Item it;
        if (!get_item(plan, t, cta_rank, &it)) continue;
        if constexpr (SKEW) {
          const int K = plan.k_blocks, H = (K + plan.gap) >> 1, T = it.tile_end - it.tile_begin;
          const int steps = (T - 1) * H + K;
          for (int st = 0; st < steps; ++st) {
            const int j0 = skew_first_tile(st, K, H);
            int j1 = st / H;
            if (j1 > T - 1) j1 = T - 1;
            const long long w1 = plan.prof ? clock64() : 0;
            mbar_wait(&full_bar[stage], phase);
            if (plan.prof) mw_full += clock64() - w1;
            tc_fence_after();
            const uint64_t da = make_smem_desc_sw128(smem_u32(smem + L::kOffA + stage * A_BYTES));
            for (int j = j0; j <= j1; ++j) {   // the older tile first: it never has to wait for a drain
              const int acc = j & 1;
              const int first = j * H;
              if (st == first) {   // this accumulator's previous tile must have been drained
                const long long w0 = plan.prof ? clock64() : 0;
                mbar_wait(&tempty_bar[acc], ((acc ? uses[1] : uses[0]) & 1) ^ 1);
                if (plan.prof) mw_tempty += clock64() - w0;
                tc_fence_after();
              }
              const uint64_t db = make_smem_desc_sw128(smem_u32(smem + L::kOffB + stage * BS_BYTES + acc * BH_BYTES));
#pragma unroll
              for (int kk = 0; kk < BK / 16; ++kk)
                umma_f16_pair(tmem_base + acc * BN, da + 2 * kk, db + 2 * kk, idesc, (st > first || kk > 0) ? 1u : 0u);
              if (st == first + K - 1) {   // every k-block seen once (cyclically): hand the tile to the epilogue
                umma_commit_pair(&tfull_bar[acc]);
                if (acc) ++uses[1]; else ++uses[0];
              }
            }
            umma_commit_pair(&empty_bar[stage]);  // frees this stage in both CTAs
            if (++stage == NST) { stage = 0; phase ^= 1; }
          }
        } else if constexpr (WIDE) {
          for (int tile = it.tile_begin; tile < it.tile_end; tile += 2) {
            const bool two = tile + 1 < it.tile_end;
            const long long w0 = plan.prof ? clock64() : 0;
            mbar_wait(&tempty_bar[0], (uses[0] & 1) ^ 1);
            if (plan.prof) mw_tempty += clock64() - w0;
            tc_fence_after();
            for (int kb = 0; kb < plan.k_blocks; ++kb) {
              const long long w1 = plan.prof ? clock64() : 0;
              mbar_wait(&full_bar[stage], phase);
              if (plan.prof) mw_full += clock64() - w1;
              tc_fence_after();
              const uint64_t da = make_smem_desc_sw128(smem_u32(smem + L::kOffA + stage * A_BYTES));
              const uint64_t db = make_smem_desc_sw128(smem_u32(smem + L::kOffB + stage * BS_BYTES));
#pragma unroll
              for (int kk = 0; kk < BK / 16; ++kk)
                umma_f16_pair(tmem_base, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
              if (two) {
                if (kb == 0) {  // the second stage drains after the first: wait as late as possible
                  const long long w2 = plan.prof ? clock64() : 0;
                  mbar_wait(&tempty_bar[1], (uses[1] & 1) ^ 1);
                  if (plan.prof) mw_tempty += clock64() - w2;
                  tc_fence_after();
                }
                const uint64_t db1 = make_smem_desc_sw128(smem_u32(smem + L::kOffB + stage * BS_BYTES + BH_BYTES));
#pragma unroll
                for (int kk = 0; kk < BK / 16; ++kk)
                  umma_f16_pair(tmem_base + BN, da + 2 * kk, db1 + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
              }
              umma_commit_pair(&empty_bar[stage]);  // frees this stage in both CTAs
              if (++stage == NST) { stage = 0; phase ^= 1; }
            }
            umma_commit_pair(&tfull_bar[0]);
            ++uses[0];
            if (two) {
              umma_commit_pair(&tfull_bar[1]);
              ++uses[1];
            }
          }
        } else {
        for (int tile = it.tile_begin; tile < it.tile_end; ++tile, ++acc_it) {
          const uint32_t acc = acc_it & 1, acc_phase = (acc_it >> 1) & 1;
          const long long w0 = plan.prof ? clock64() : 0;
          mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
          if (plan.prof) mw_tempty += clock64() - w0;
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * BN;
          for (int kb = 0; kb < plan.k_blocks; ++kb) {
            const long long w1 = plan.prof ? clock64() : 0;
            mbar_wait(&full_bar[stage], phase);
            if (plan.prof) mw_full += clock64() - w1;
            tc_fence_after();
            const uint64_t da = make_smem_desc_sw128(smem_u32(smem + L::kOffA + stage * A_BYTES));
            const uint64_t db = make_smem_desc_sw128(smem_u32(smem + L::kOffB + stage * BS_BYTES));
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk)
              umma_f16_pair(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
            umma_commit_pair(&empty_bar[stage]);  // frees this stage in both CTAs
            if (++stage == NST) { stage = 0; phase ^= 1; }
          }
          umma_commit_pair(&tfull_bar[acc]);      // both CTAs' epilogues may read their rows
        }
        }
      }
      if (plan.prof) {
        plan.prof[blockIdx.x * 16 + 0] = clock64() - m_t0;
        plan.prof[blockIdx.x * 16 + 1] = mw_full;
        plan.prof[blockIdx.x * 16 + 2] = mw_tempty;
      }
    }
  } else if (warp == 6) {
    if (ga.world > 1) gather_warp(plan, ga, lane);
    if (dg.world > 1) data_gather_warp(plan, dg, lane);
  } else {
    const int half = HALVES && warp > 6 ? 1 : 0;
    const int off_list = L::kOffList + half * BM * CAP * 8;
    if (cta_rank == 0)
      epilogue_loop<false>(plan, smem, off_list, L::kOffXn, tmem_base, tfull_bar, tempty_bar, cluster_id,
                           num_clusters, cta_rank, total_items, xnorm2, cand, half, HALVES ? 2 : 1);
    else
      epilogue_loop<true>(plan, smem, off_list, L::kOffXn, tmem_base, tfull_bar, tempty_bar, cluster_id,
                          num_clusters, cta_rank, total_items, xnorm2, cand, half, HALVES ? 2 : 1);
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
  // fused row all-gather: once every CTA's communication warp is through, tell the row's peers
  // that their parts may be overwritten ("done reading")
  signal_peers_when_grid_done(dg.done);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void *p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess) return nullptr;
  if (qres != cudaDriverEntryPointSuccess) return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

}  // namespace

int make_fp16_tile_map(CUtensorMap *map, const __half *base, long long rows, int a_pad, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  DMLP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(a_pad), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(a_pad) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(BK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half *>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DMLP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (code " + std::to_string(int(r)) + ")");
  return 0;
}

namespace {
TopkPlan make_single_plan(int q_rows, int n, int a_pad, int kp, int num_sms, int kmax);

// Sets of epilogue warps (= lists per query and column chunk).  Short rows -- one or two k-blocks
// per tile -- are bound by the epilogue, so the CTA-pair kernel runs two sets there (MODE 3).
int plan_halves(int q_rows, int a_pad) {
  bool pair = (q_rows + BM - 1) / BM >= 2;
  if (const char *e = getenv("DMLP_TOPK_PAIR")) pair = pair && atoi(e) != 0;
  bool wide = a_pad / BK >= 64;
  if (const char *e = getenv("DMLP_TOPK_WIDE")) wide = atoi(e) != 0;
  // Measured on B200 (70000 x 200000, profiles/ab_append_r2.jsonl): A = 64 15.0 ms with two sets against
  // 16.1 with one; A = 512 9.9 against 8.6, A = 1024 17.7 against 14.3.  The gain is small because the
  // epilogue's time is mostly in the appends, and two half-width lists per row take 1.8 x the appends
  // of one (every list warms its threshold up by itself); with more than one k-block per tile the
  // two-stage ring of this mode costs more than the second set returns.  So: one k-block only.
  int halves = a_pad / BK <= 1 ? 2 : 1;
  if (const char *e = getenv("DMLP_TOPK_HALVES")) halves = atoi(e) == 2 ? 2 : 1;
  return (pair && !wide) ? halves : 1;
}
}

// Large k: every (query, chunk) list keeps 32 keys and the certified filter wants one list per 8 of
// k (the ~1.6 k survivors then fill a list to 13 of 32 on average: a full list, which would flag the
// query, stays a rare event).  One launch has at most 32 chunks, so for k > 256 the data rows are cut
// into `parts` ranges of at least 32 tiles and every range gets its own launch and its own lists of
// the shared candidate buffer.  Needs n >= 8192 * parts (about 32 k); below that the caller's exact
// path is the better tool anyway (k is then more than 3 % of the data).
TopkPlan make_topk_plan(int q_rows, int n, int a_pad, int kp, int num_sms, int kmax) {
  const int c_need = (kmax + 7) / 8;   // lists
  const int per_launch = 32 * plan_halves(q_rows, a_pad);
  const int parts = (c_need + per_launch - 1) / per_launch;
  const int n_tiles = (n + BN - 1) / BN;
  if (parts > 1 && n_tiles >= 32 * parts) {
    const int part_kmax = (kmax + parts - 1) / parts;
    const int tiles0 = (n_tiles + parts - 1) / parts;
    TopkPlan p = make_single_plan(q_rows, tiles0 * BN, a_pad, kp, num_sms, part_kmax);
    p.n = n;
    p.n_tiles = n_tiles;
    p.parts = parts;
    p.part_chunks = p.lists;
    p.part_kmax = part_kmax;
    p.lists = parts * p.part_chunks;
    p.cand_chunks = p.lists;
    return p;
  }
  return make_single_plan(q_rows, n, a_pad, kp, num_sms, kmax);
}

namespace {
TopkPlan make_single_plan(int q_rows, int n, int a_pad, int kp, int num_sms, int kmax) {
  TopkPlan p;
  p.parts = 1;
  p.part_chunks = 0;
  p.part_kmax = kmax;
  p.num_sms = num_sms;
  p.q_rows = q_rows;
  p.n = n;
  p.a_pad = a_pad;
  p.kp = kp;
  p.m_blocks = (q_rows + BM - 1) / BM;
  p.n_tiles = (n + BN - 1) / BN;
  p.k_blocks = a_pad / BK;
  if (num_sms < 1) num_sms = 1;
  long long budget_mb = 160;
  int cluster = 4;
  p.hint_q = 2;  // evict-last
  p.hint_x = 0;
  if (const char *e = getenv("DMLP_TOPK_L2_MB")) budget_mb = atoll(e) > 0 ? atoll(e) : budget_mb;
  if (const char *e = getenv("DMLP_TOPK_HINT_Q")) p.hint_q = atoi(e);
  if (const char *e = getenv("DMLP_TOPK_HINT_X")) p.hint_x = atoi(e);
  if (const char *e = getenv("DMLP_TOPK_CLUSTER")) cluster = atoi(e);
  p.pair = 1;  // CTA-pair MMA (cta_group::2) unless there is a single row block
  if (const char *e = getenv("DMLP_TOPK_PAIR")) p.pair = atoi(e) != 0;
  if (p.m_blocks < 2) p.pair = 0;
  if (p.pair) cluster = 2;
  // wide tiles pay when the k-loop is long enough to amortise draining both accumulators
  // with the MMA idle (measured: A = 8192 11 % faster, A = 1024 27 % slower)
  p.wide = p.k_blocks >= 64;
  p.prof = nullptr;
  if (const char *e = getenv("DMLP_TOPK_WIDE")) p.wide = atoi(e) != 0;
  if (!p.pair) p.wide = 0;
  // skewed wide schedule (opt-in): steps an accumulator rests between two tiles (its drain; the
  // other accumulator keeps the tensor pipe busy meanwhile).  DMLP_TOPK_GAP=<steps>; the default
  // -1 is the lock-step wide tile (both accumulators fill side by side and drain while the MMA waits)
  p.gate_cnt = nullptr;
  p.gate_epoch = 0;
  p.gap = -1;   // measured on B200: the skewed schedule is slower (65.9 vs 57 ms on the headline shape), see DESIGN.md
  if (const char *e = getenv("DMLP_TOPK_GAP")) p.gap = atoi(e);
  if (p.gap > p.k_blocks - 2) p.gap = p.k_blocks - 2;
  if (p.gap >= 0 && ((p.k_blocks + p.gap) & 1)) ++p.gap;   // the half period must be whole
  if (p.gap > p.k_blocks - 2) p.gap -= 2;
  if (p.gap < 0 || !p.wide) p.gap = -1;
  // -1 = automatic: dynamic only where the launch runs without re-alignment (short rows).  Measured on
  // B200 (profiles/ab_dyn_r2.jsonl): +2.6 % at A = 64, +-0 at A = 512, -14..-17 % at A = 8192, where the
  // static waves are balanced anyway and the re-alignment of co-resident clusters is worth 23 %.
  p.dyn = -1;
  if (const char *e = getenv("DMLP_TOPK_DYN")) p.dyn = atoi(e) != 0;
  p.dyn_counter = nullptr;
  p.dyn_slots = nullptr;
  p.dyn_cap = 0;
  p.rot = 0;
  p.cand_chunks = 0;  // filled in once the chunk count is known
  p.chunk_base = 0;
  p.row_base = 0;
  // re-align the clusters of a chunk every `sync` data tiles: about once per 128 k-blocks of
  // work (a tile of the headline shape); problems with short rows are epilogue-bound and
  // their tiles are small, so they skip it.  DMLP_TOPK_SYNC=<tiles>, 0 = off
  p.sync = p.k_blocks >= 16 ? (128 + p.k_blocks - 1) / p.k_blocks : 0;
  p.sync_cnt = nullptr;
  if (const char *e = getenv("DMLP_TOPK_SYNC")) p.sync = atoi(e);
  if (p.sync < 0) p.sync = 0;

  // Cluster: CTAs that sweep the same data tiles in lock step and share each
  // tile through TMA multicast.  Never wider than the number of row blocks.
  if (cluster != 1 && cluster != 2 && cluster != 4 && cluster != 8) cluster = 4;
  while (cluster > 1 && cluster > p.m_blocks) cluster >>= 1;
  p.cluster = cluster;
  p.mc_blocks = (p.m_blocks + cluster - 1) / cluster;
  // Clusters that fit the chip at once (GPC boundaries strand a few SMs: 148 SMs
  // hold 74 pairs, 33 quads or 16 octets).
  int nclusters = cluster == 1 ? num_sms : cluster == 2 ? num_sms / 2 : cluster == 4 ? (num_sms * 33) / 148
                                                                                      : (num_sms * 16) / 148;
  if (nclusters < 1) nclusters = 1;
  // Split the clusters into `chunks` column ranges x `group_mc` cluster rows; one wave
  // processes one group (group_mc x chunks clusters), the row blocks are covered group
  // by group.  Wanted: (1) every wave fills the chip -- with few row blocks that takes
  // more chunks; (2) many cluster rows per chunk, because they sweep the same data tiles
  // together and share them through L2 (measured fastest at two chunks on the headline
  // shape); (3) the wave's query tiles within reach of L2; (4) at least two chunks when
  // possible: each contributes 32 keys per query, the head-room the certified filter
  // needs above k.
  long long tile_bytes = static_cast<long long>(BM) * a_pad * 2 * cluster;
  long long budget = budget_mb << 20;
  // the certified survivor set is about 1.6 k per query, spread over the chunks, and a
  // chunk can contribute at most 32: one chunk per 8 of k keeps every list far from full
  p.halves = plan_halves(q_rows, a_pad);
  const int c_need = ((kmax + 7) / 8 + p.halves - 1) / p.halves;
  int c_min = 2 > c_need ? 2 : c_need;
  if (c_min > p.n_tiles) c_min = p.n_tiles > 0 ? p.n_tiles : 1;
  int c_max = p.n_tiles < 32 ? (p.n_tiles > 0 ? p.n_tiles : 1) : 32;
  if (c_max > nclusters) c_max = nclusters;
  if (c_min > c_max) c_min = c_max;
  // Few, long chunks are best (every (row, chunk) sweep starts with an open threshold and
  // pays a warm-up of list compactions), so: the fewest chunks allowed, the most cluster
  // rows per wave that L2 takes, row blocks spread evenly over the waves, and whatever
  // clusters that leaves idle go to extra chunks.
  int g_max = nclusters / c_min;
  while (g_max > 1 && g_max * tile_bytes > budget) --g_max;
  if (g_max < 1) g_max = 1;
  const int groups = (p.mc_blocks + g_max - 1) / g_max;
  int gmc = groups > 0 ? (p.mc_blocks + groups - 1) / groups : 1;
  if (gmc < 1) gmc = 1;
  int c = nclusters / gmc;
  if (c < c_min) c = c_min;
  if (c > c_max) c = c_max;
  if (const char *e = getenv("DMLP_TOPK_CHUNKS")) c = atoi(e) > 0 ? atoi(e) : c;
  if (const char *e = getenv("DMLP_TOPK_GMC")) gmc = atoi(e) > 0 ? atoi(e) : gmc;
  if (getenv("DMLP_TOPK_CHUNKS") && !getenv("DMLP_TOPK_GMC")) gmc = nclusters / c > 0 ? nclusters / c : 1;
  if (gmc > p.mc_blocks) gmc = p.mc_blocks > 0 ? p.mc_blocks : 1;
  if (c > p.n_tiles) c = p.n_tiles > 0 ? p.n_tiles : 1;
  if (c < 1) c = 1;
  p.group_mc = gmc;
  p.group_m = gmc * cluster;
  p.chunks = c;
  p.cand_chunks = c;
  long long items = static_cast<long long>((p.mc_blocks + gmc - 1) / gmc) * gmc * c;
  int use = static_cast<int>(items < nclusters ? (items > 0 ? items : 1) : nclusters);
  // one wave = one whole group (gmc x c clusters), so that clusters sharing a chunk run side by side
  if (p.pair && static_cast<long long>(gmc) * c <= use) use = gmc * c;
  p.grid = use * cluster;
  // short rows (one or two k-blocks per tile) are bound by the epilogue: two sets of epilogue warps
  if (!p.pair || p.wide) p.halves = 1;
  p.lists = p.chunks * p.halves;
  p.cand_chunks = p.lists;
  p.part_chunks = p.lists;
  return p;
}
}  // namespace

namespace {
// Re-alignment counters: one file-scope array per device, handed out in slices.  Every launch
// takes the next `need` words (a host-side cursor that wraps), so launches that overlap on
// different streams count on different words; only more than kSyncCounters words' worth of
// launches in flight at once could meet on the same slice.
constexpr int kSyncCounters = 16384;
__device__ unsigned g_sync_counters[kSyncCounters];
std::atomic<unsigned long long> g_sync_cursor{0};
// Dynamic scheduler scratch: counter + per-cluster slots, handed out in slices like the counters above.
constexpr int kDynWords = 1 << 20;
__device__ unsigned g_dyn_words[kDynWords];
std::atomic<unsigned long long> g_dyn_cursor{0};
constexpr int kProfCtas = 256;
__device__ unsigned long long g_prof[kProfCtas * 16];
int g_prof_ctas = 0;
}  // namespace

int enumerate_topk_items(const TopkPlan &plan, int *out, int max_items) {
  const int num_clusters = plan.grid / plan.cluster;
  const int per_group = plan.group_mc * plan.chunks;
  const int total_items = ((plan.mc_blocks + plan.group_mc - 1) / plan.group_mc) * per_group;
  int n = 0;
  for (int cluster_id = 0; cluster_id < num_clusters; ++cluster_id)
    for (int t = cluster_id; t < total_items; t += num_clusters)
      for (int rank = 0; rank < plan.cluster; ++rank) {
        Item it;
        if (!get_item(plan, t, rank, &it)) continue;
        if (n < max_items) {
          int *o = out + static_cast<long long>(n) * 8;
          o[0] = cluster_id; o[1] = rank; o[2] = it.m; o[3] = it.chunk;
          o[4] = it.tile_begin; o[5] = it.tile_end; o[6] = it.group; o[7] = it.group_rows;
        }
        ++n;
      }
  return n;
}

int dist_topk_sync_timeouts(unsigned *out) {
  DMLP_CUDA_OK(cudaMemcpyFromSymbol(out, g_sync_timeouts, sizeof(unsigned)));
  return 0;
}

int dist_topk_prof_read(unsigned long long *host_out, int max_ctas) {
  const int n = g_prof_ctas < max_ctas ? g_prof_ctas : max_ctas;
  if (n <= 0) return 0;
  DMLP_CUDA_OK(cudaDeviceSynchronize());
  DMLP_CUDA_OK(cudaMemcpyFromSymbol(host_out, g_prof, sizeof(unsigned long long) * 16 * n));
  return n;
}

int launch_dist_topk(const __half *qh, const __half *xh, const float *xnorm2, const TopkPlan &plan_in,
                     unsigned long long *cand, const GatherArgs *gather, cudaStream_t s,
                     const DataGatherArgs *data_gather) {
  TopkPlan plan = plan_in;
  if (plan.q_rows <= 0 || plan.n <= 0) return 0;
  if (plan.parts > 1) {
    DMLP_REQUIRE((gather == nullptr || gather->world <= 1) && (data_gather == nullptr || data_gather->world <= 1),
                 "dist_topk: the in-kernel gathers need a single launch (k <= 256)");
    for (int p = 0; p < plan.parts; ++p) {
      const int t0 = static_cast<int>(static_cast<long long>(p) * plan.n_tiles / plan.parts);
      const int t1 = static_cast<int>(static_cast<long long>(p + 1) * plan.n_tiles / plan.parts);
      const int r0 = t0 * BN, r1 = t1 * BN < plan.n ? t1 * BN : plan.n;
      TopkPlan sub = make_single_plan(plan.q_rows, r1 - r0, plan.a_pad, plan.kp, plan.num_sms, plan.part_kmax);
      DMLP_REQUIRE(sub.lists == plan.part_chunks, "dist_topk: the parts of a large-k launch must agree on their list count");
      sub.cand_chunks = plan.cand_chunks;
      sub.chunk_base = plan.chunk_base + p * plan.part_chunks;
      sub.row_base = plan.row_base + r0;
      if (int rc = launch_dist_topk(qh, xh + static_cast<size_t>(r0) * plan.a_pad, xnorm2 + r0, sub, cand, nullptr, s,
                                    nullptr))
        return rc;
    }
    return 0;
  }
  {
    // per-launch tile-alignment counters (internal scratch, one set per device)
    const int groups = (plan.mc_blocks + plan.group_mc - 1) / plan.group_mc;
    const long long need = static_cast<long long>(groups) * plan.chunks;
    // the alignment wait needs every cluster of a group on the chip at once: only when the
    // grid is a single wave of whole groups
    const bool whole_wave = plan.pair && plan.group_mc * plan.chunks * plan.cluster == plan.grid;
    if (plan.sync && whole_wave && need <= kSyncCounters) {
      void *p = nullptr;
      DMLP_CUDA_OK(cudaGetSymbolAddress(&p, g_sync_counters));
      // claim a slice [off, off + need) that does not straddle the end of the array
      unsigned long long cur = g_sync_cursor.load(), off, next;
      do {
        off = cur % kSyncCounters;
        if (off + need > kSyncCounters) off = 0;
        next = off + need;
      } while (!g_sync_cursor.compare_exchange_weak(cur, next));
      unsigned *slice = static_cast<unsigned *>(p) + off;
      DMLP_CUDA_OK(cudaMemsetAsync(slice, 0, sizeof(unsigned) * need, s));
      plan.sync_cnt = slice;
    } else {
      plan.sync = 0;
    }
  }
  plan.dyn_counter = plan.dyn_slots = nullptr;
  if ((plan.dyn > 0 || (plan.dyn < 0 && plan.sync == 0)) && plan.pair && (gather == nullptr || gather->world <= 1) &&
      (data_gather == nullptr || data_gather->world <= 1)) {
    // the gather warps of the fused multi-GPU launches walk the static item list: those stay static
    const int groups = (plan.mc_blocks + plan.group_mc - 1) / plan.group_mc;
    const long long total_items = static_cast<long long>(groups) * plan.group_mc * plan.chunks;
    const long long cap = total_items + 1;     // every grab of one cluster, plus its end marker
    const long long need = 1 + cap * (plan.grid / plan.cluster);
    if (need <= kDynWords) {
      void *p = nullptr;
      DMLP_CUDA_OK(cudaGetSymbolAddress(&p, g_dyn_words));
      unsigned long long cur = g_dyn_cursor.load(), off, next;
      do {
        off = cur % kDynWords;
        if (off + need > kDynWords) off = 0;
        next = off + need;
      } while (!g_dyn_cursor.compare_exchange_weak(cur, next));
      unsigned *slice = static_cast<unsigned *>(p) + off;
      DMLP_CUDA_OK(cudaMemsetAsync(slice, 0, sizeof(unsigned) * need, s));
      plan.dyn_counter = slice;
      plan.dyn_slots = slice + 1;
      plan.dyn_cap = static_cast<int>(cap);
      plan.sync = 0;   // a group's clusters no longer start its items together
      plan.sync_cnt = nullptr;
    }
  }
  plan.prof = nullptr;
  if (getenv("DMLP_TOPK_PROF") && plan.pair && plan.grid <= kProfCtas) {
    void *p = nullptr;
    DMLP_CUDA_OK(cudaGetSymbolAddress(&p, g_prof));
    DMLP_CUDA_OK(cudaMemsetAsync(p, 0, sizeof(unsigned long long) * 16 * plan.grid, s));
    plan.prof = static_cast<unsigned long long *>(p);
    g_prof_ctas = plan.grid;
  }
  if (!plan.pair && plan.cluster > 1) {
    // TMA multicast loads here read only this GPU's memory
    int dev = -1;
    DMLP_CUDA_OK(cudaGetDevice(&dev));
    cudaPointerAttributes at_x;
    DMLP_CUDA_OK(cudaPointerGetAttributes(&at_x, xh));
    DMLP_REQUIRE(at_x.type == cudaMemoryTypeDevice && at_x.device == dev,
                 "dist_topk: the multicast tile loads need the data operand in this GPU's own memory");
  }
  DMLP_REQUIRE(plan.kp == KP, "dist_topk: this build keeps 32 keys per (query, chunk)");
  DMLP_REQUIRE(plan.a_pad % BK == 0 && plan.a_pad > 0, "dist_topk: a_pad must be a positive multiple of 64");
  CUtensorMap tm_q, tm_x;
  if (int rc = make_fp16_tile_map(&tm_q, qh, plan.q_rows, plan.a_pad, BM)) return rc;
  if (int rc = make_fp16_tile_map(&tm_x, xh, plan.n, plan.a_pad, plan.pair ? BN / 2 : BN / plan.cluster)) return rc;
  if (!plan.pair) plan.wide = 0;
  if (!plan.pair || plan.wide) plan.halves = 1;
  const bool halves = plan.halves == 2;
  const int smem_bytes =
      plan.pair ? (plan.wide ? Layout1::kBytes : (halves ? Layout3::kBytes : Layout2::kBytes)) : Layout1::kBytes;
  if (!plan.wide) plan.gap = -1;
  if (halves)
    DMLP_CUDA_OK(cudaFuncSetAttribute(dist_topk_pair_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  else if (plan.pair && plan.wide && plan.gap >= 0)
    DMLP_CUDA_OK(cudaFuncSetAttribute(dist_topk_pair_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  else if (plan.pair && plan.wide)
    DMLP_CUDA_OK(cudaFuncSetAttribute(dist_topk_pair_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  else if (plan.pair)
    DMLP_CUDA_OK(cudaFuncSetAttribute(dist_topk_pair_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  else
    DMLP_CUDA_OK(cudaFuncSetAttribute(dist_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(plan.grid);
  cfg.blockDim = dim3(plan.pair ? (halves ? THREADS3 : THREADS2) : THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  GatherArgs ga;
  if (gather != nullptr) ga = *gather;
  DMLP_REQUIRE(ga.world == 1 || plan.pair, "dist_topk: the fused operand gather needs the CTA-pair kernel");
  DMLP_REQUIRE(ga.world <= kMaxPeers, "dist_topk: at most 8 ranks share one gather");
  DataGatherArgs dg;
  if (data_gather != nullptr) dg = *data_gather;
  plan.gate_cnt = dg.world > 1 ? dg.tile_count : nullptr;
  plan.gate_epoch = dg.epoch;
  DMLP_REQUIRE(dg.world == 1 || plan.pair, "dist_topk: the fused data gather needs the CTA-pair kernel");
  DMLP_REQUIRE(dg.world <= kMaxPeers, "dist_topk: at most 8 ranks share one gather");
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = plan.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (halves)
    DMLP_CUDA_OK(cudaLaunchKernelEx(&cfg, dist_topk_pair_kernel<3>, tm_q, tm_x, xnorm2, plan, cand, ga, dg));
  else if (plan.pair && plan.wide && plan.gap >= 0)
    DMLP_CUDA_OK(cudaLaunchKernelEx(&cfg, dist_topk_pair_kernel<2>, tm_q, tm_x, xnorm2, plan, cand, ga, dg));
  else if (plan.pair && plan.wide)
    DMLP_CUDA_OK(cudaLaunchKernelEx(&cfg, dist_topk_pair_kernel<1>, tm_q, tm_x, xnorm2, plan, cand, ga, dg));
  else if (plan.pair)
    DMLP_CUDA_OK(cudaLaunchKernelEx(&cfg, dist_topk_pair_kernel<0>, tm_q, tm_x, xnorm2, plan, cand, ga, dg));
  else
    DMLP_CUDA_OK(cudaLaunchKernelEx(&cfg, dist_topk_kernel, tm_q, tm_x, xnorm2, plan, cand));
  return 0;
}

}  // namespace k
}  // namespace dmlp
