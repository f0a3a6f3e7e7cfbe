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

constexpr int OFF_A = 0;
constexpr int OFF_B = OFF_A + STAGES * A_BYTES;
constexpr int OFF_LIST = OFF_B + STAGES * B_BYTES;
constexpr int OFF_XN = OFF_LIST + BM * CAP * 8;
constexpr int OFF_BAR = OFF_XN + 2 * BN * 4;
constexpr int SMEM_BYTES = OFF_BAR + 128 + 1024;  // + alignment slack

struct Item {
  int m;
  int chunk;
  int tile_begin, tile_end;
};

__device__ __forceinline__ bool get_item(const TopkPlan &p, int t, Item *it) {
  const int per_group = p.group_m * p.chunks;
  const int g = t / per_group;
  const int rem = t - g * per_group;
  const int m0 = g * p.group_m;
  const int gm = min(p.group_m, p.m_blocks - m0);
  if (gm <= 0 || rem >= gm * p.chunks) return false;
  const int c = rem / gm;
  it->m = m0 + (rem - c * gm);
  it->chunk = c;
  it->tile_begin = static_cast<int>(static_cast<long long>(c) * p.n_tiles / p.chunks);
  it->tile_end = static_cast<int>(static_cast<long long>(c + 1) * p.n_tiles / p.chunks);
  return true;
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
// `emit` the sorted keys are also written to global memory.
__device__ __forceinline__ void compact_row(unsigned long long *row_keys, int rot, int r, int lane, int &cnt,
                                            float &tau, unsigned long long *emit) {
  const int rcnt = __shfl_sync(0xFFFFFFFFu, cnt, r);
  unsigned long long a = (lane < rcnt) ? row_keys[(lane + rot) & (CAP - 1)] : kKeyInvalid;
  unsigned long long b = (lane + 32 < rcnt) ? row_keys[(lane + 32 + rot) & (CAP - 1)] : kKeyInvalid;
  warp_sort64(a, b, lane);
  row_keys[(lane + rot) & (CAP - 1)] = a;
  const unsigned long long last = __shfl_sync(0xFFFFFFFFu, a, KP - 1);
  if (lane == r) {
    cnt = rcnt < KP ? rcnt : KP;
    tau = (rcnt >= KP) ? key_score(last) : std::numeric_limits<float>::infinity();
  }
  if (emit != nullptr) emit[lane] = a;
  __syncwarp();
}

__global__ void __launch_bounds__(THREADS, 1)
dist_topk_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_x,
                 const float *__restrict__ xnorm2, const TopkPlan plan, unsigned long long *__restrict__ cand) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + OFF_BAR);
  uint64_t *full_bar = bars;                  // [STAGES]
  uint64_t *empty_bar = bars + STAGES;        // [STAGES]
  uint64_t *tfull_bar = bars + 2 * STAGES;    // [2]
  uint64_t *tempty_bar = bars + 2 * STAGES + 2;  // [2]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_x);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
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
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int per_group = plan.group_m * plan.chunks;
  const int total_items = ((plan.m_blocks + plan.group_m - 1) / plan.group_m) * per_group;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_items; t += gridDim.x) {
        Item it;
        if (!get_item(plan, t, &it)) continue;
        for (int tile = it.tile_begin; tile < it.tile_end; ++tile) {
          for (int kb = 0; kb < plan.k_blocks; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
            tma_load_2d(smem + OFF_A + stage * A_BYTES, &tm_q, &full_bar[stage], kb * BK, it.m * BM);
            tma_load_2d(smem + OFF_B + stage * B_BYTES, &tm_x, &full_bar[stage], kb * BK, tile * BN);
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
      for (int t = blockIdx.x; t < total_items; t += gridDim.x) {
        Item it;
        if (!get_item(plan, t, &it)) continue;
        for (int tile = it.tile_begin; tile < it.tile_end; ++tile, ++acc_it) {
          const uint32_t acc = acc_it & 1, acc_phase = (acc_it >> 1) & 1;
          mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * BN;
          for (int kb = 0; kb < plan.k_blocks; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint64_t da = make_smem_desc_sw128(smem_u32(smem + OFF_A + stage * A_BYTES));
            const uint64_t db = make_smem_desc_sw128(smem_u32(smem + OFF_B + stage * B_BYTES));
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk)
              umma_f16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0 ? 1u : 0u);
            umma_commit(&empty_bar[stage]);  // frees the smem stage once these MMAs retire
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          umma_commit(&tfull_bar[acc]);      // accumulator complete -> epilogue
        }
      }
    }
  } else {
    // -------------------------------- epilogue --------------------------------
    const int quarter = warp & 3;              // tensor-memory lane quarter this warp may read
    const int row_in_tile = quarter * 32 + lane;
    unsigned long long *lists = reinterpret_cast<unsigned long long *>(smem + OFF_LIST);
    unsigned long long *my_keys = lists + row_in_tile * CAP;
    float *xn_all = reinterpret_cast<float *>(smem + OFF_XN);
    const int e_tid = threadIdx.x - 64;
    const float inf = std::numeric_limits<float>::infinity();
    int cnt = 0;
    float tau = inf;
    uint32_t acc_it = 0;

    for (int t = blockIdx.x; t < total_items; t += gridDim.x) {
      Item it;
      if (!get_item(plan, t, &it)) continue;
      for (int tile = it.tile_begin; tile < it.tile_end; ++tile, ++acc_it) {
        const uint32_t acc = acc_it & 1, acc_phase = (acc_it >> 1) & 1;
        float *xn = xn_all + acc * BN;
        for (int i = e_tid; i < BN; i += EPI_THREADS) {
          const int col = tile * BN + i;
          xn[i] = col < plan.n ? __ldg(xnorm2 + col) : inf;
        }
        named_barrier_sync(1, EPI_THREADS);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
          // make room for up to 32 appends per row
          __syncwarp();
          unsigned need = __ballot_sync(0xFFFFFFFFu, cnt > CAP - 32);
          while (need) {
            const int r = __ffs(need) - 1;
            need &= need - 1;
            compact_row(lists + (quarter * 32 + r) * CAP, quarter * 32 + r, r, lane, cnt, tau, nullptr);
          }
          uint32_t v[32];
          tmem_ld_32x32(t_row + c0, v);
          tmem_ld_wait();
          float s[32];
          float mn = inf;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            s[j] = fmaf(-2.0f, __uint_as_float(v[j]), xn[c0 + j]);
            mn = fminf(mn, s[j]);
          }
          if (mn < tau) {
            const int id0 = tile * BN + c0;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (s[j] < tau) {
                my_keys[(cnt + row_in_tile) & (CAP - 1)] = make_key(s[j], id0 + j);
                ++cnt;
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      }
      // flush this (row block, chunk): sorted kp keys per row
      __syncwarp();
      for (int r = 0; r < 32; ++r) {
        const long long qrow = static_cast<long long>(it.m) * BM + quarter * 32 + r;
        unsigned long long *out =
            qrow < plan.q_rows ? cand + (qrow * plan.chunks + it.chunk) * KP : nullptr;
        compact_row(lists + (quarter * 32 + r) * CAP, quarter * 32 + r, r, lane, cnt, tau, out);
      }
      cnt = 0;
      tau = inf;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
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

TopkPlan make_topk_plan(int q_rows, int n, int a_pad, int kp, int num_sms) {
  TopkPlan p;
  p.q_rows = q_rows;
  p.n = n;
  p.a_pad = a_pad;
  p.kp = kp;
  p.m_blocks = (q_rows + BM - 1) / BM;
  p.n_tiles = (n + BN - 1) / BN;
  p.k_blocks = a_pad / BK;
  if (num_sms < 1) num_sms = 1;
  // Row blocks per wave: keep their query tiles (128 x a_pad fp16 each) within
  // a 32 MB slice of L2 so only the data tiles stream from HBM.
  long long tile_bytes = static_cast<long long>(BM) * a_pad * 2;
  long long fit = (32LL << 20) / (tile_bytes > 0 ? tile_bytes : 1);
  int gm = static_cast<int>(fit < 1 ? 1 : (fit > num_sms ? num_sms : fit));
  if (gm > p.m_blocks) gm = p.m_blocks > 0 ? p.m_blocks : 1;
  p.group_m = gm;
  int c = num_sms / gm;
  if (c < 1) c = 1;
  if (c > p.n_tiles) c = p.n_tiles > 0 ? p.n_tiles : 1;
  p.chunks = c;
  long long items = static_cast<long long>((p.m_blocks + gm - 1) / gm) * gm * c;
  p.grid = static_cast<int>(items < num_sms ? (items > 0 ? items : 1) : num_sms);
  return p;
}

int launch_dist_topk(const __half *qh, const __half *xh, const float *xnorm2, const TopkPlan &plan,
                     unsigned long long *cand, cudaStream_t s) {
  if (plan.q_rows <= 0 || plan.n <= 0) return 0;
  DMLP_REQUIRE(plan.kp == KP, "dist_topk: this build keeps 32 keys per (query, chunk)");
  DMLP_REQUIRE(plan.a_pad % BK == 0 && plan.a_pad > 0, "dist_topk: a_pad must be a positive multiple of 64");
  CUtensorMap tm_q, tm_x;
  if (int rc = make_fp16_tile_map(&tm_q, qh, plan.q_rows, plan.a_pad, BM)) return rc;
  if (int rc = make_fp16_tile_map(&tm_x, xh, plan.n, plan.a_pad, BN)) return rc;
  DMLP_CUDA_OK(cudaFuncSetAttribute(dist_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  dist_topk_kernel<<<plan.grid, THREADS, SMEM_BYTES, s>>>(tm_q, tm_x, xnorm2, plan, cand);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
