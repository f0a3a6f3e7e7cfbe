// Shared device helpers for the sm_100a kernels: PTX wrappers for mbarrier,
// TMA and tcgen05, ordered keys, and error plumbing for the C ABI.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda.h>

#include <cstdint>
#include <cstdio>
#include <string>

namespace dmlp {
namespace k {

// ---------------------------------------------------------------------------
// host-side error plumbing (thread-local message, negative return codes)
// ---------------------------------------------------------------------------
void set_error(const std::string &msg);
const char *get_error();

#define DMLP_CUDA_OK(expr)                                                            \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) {                                                          \
      ::dmlp::k::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));       \
      return -1;                                                                      \
    }                                                                                 \
  } while (0)

#define DMLP_REQUIRE(cond, msg)                                                       \
  do {                                                                                \
    if (!(cond)) {                                                                    \
      ::dmlp::k::set_error(std::string(msg));                                         \
      return -2;                                                                      \
    }                                                                                 \
  } while (0)

// ---------------------------------------------------------------------------
// ordered keys: (approximate score, row id) packed so that unsigned 64-bit
// comparison orders by score ascending, then id ascending
// ---------------------------------------------------------------------------
constexpr unsigned long long kKeyInvalid = 0xFFFFFFFFFFFFFFFFULL;

__host__ __device__ inline uint32_t float_to_ordered(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f);
#else
  union { float f; uint32_t u; } cv; cv.f = f; uint32_t b = cv.u;
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float ordered_to_float(uint32_t o) {
  uint32_t b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  union { float f; uint32_t u; } cv; cv.u = b; return cv.f;
#endif
}
__host__ __device__ inline unsigned long long make_key(float score, int id) {
  return (static_cast<unsigned long long>(float_to_ordered(score)) << 32) | static_cast<uint32_t>(id);
}
__host__ __device__ inline float key_score(unsigned long long key) {
  return ordered_to_float(static_cast<uint32_t>(key >> 32));
}
__host__ __device__ inline int key_id(unsigned long long key) {
  return static_cast<int>(static_cast<uint32_t>(key & 0xFFFFFFFFULL));
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------
// "decimal6": a value of the text grammar ("%.6f", generate_input.py:15,20) is i / 1e6 for an
// integer i; the double a parser produces for it is the correctly rounded quotient.  One multiply
// by RN(1e-6) and Markstein's correction step (the exact remainder folded back in) give exactly
// that double with three fp64 operations instead of a division -- checked exhaustively against
// IEEE division for all 2^32 values of a 32-bit i (tests/test_cpu_oracle.py runs a sample).
// ---------------------------------------------------------------------------
__device__ __forceinline__ double dec6_to_double(int i) {
  // (double)i without the conversion pipe: 2^52 + 2^31 + i has the bits {0x43300000, i ^ 0x80000000};
  // subtracting 2^52 + 2^31 is exact
  const double x = __dsub_rn(__hiloint2double(0x43300000, i ^ static_cast<int>(0x80000000u)), 4503601774854144.0);
  const double v0 = __dmul_rn(x, 1e-6);
  const double r = __fma_rn(-1e6, v0, x);
  return __fma_rn(r, 1e-6, v0);
}

// ---------------------------------------------------------------------------
// PTX: shared-memory addressing, mbarrier
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ---------------------------------------------------------------------------
// PTX: TMA (cp.async.bulk.tensor) 2-D tile load into swizzled shared memory
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
// Multicast tile load: the tile lands at the same shared-memory offset of every CTA
// in `cta_mask`, and each of those CTAs' barriers (same offset) receives the bytes.
__device__ __forceinline__ void tma_load_2d_mcast(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0,
                                                  int c1, uint16_t cta_mask, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      ".L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5, %6;"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
      "h"(cta_mask), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// L2 eviction-priority policies for TMA loads (createpolicy encodings).
constexpr uint64_t kL2EvictNormal = 0x1000000000000000ULL;
constexpr uint64_t kL2EvictFirst = 0x12F0000000000000ULL;
constexpr uint64_t kL2EvictLast = 0x14F0000000000000ULL;
__device__ __forceinline__ void tma_load_2d_hint(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0,
                                                 int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
      "l"(policy)
      : "memory");
}

// ---------------------------------------------------------------------------
// PTX: tcgen05 (5th-generation tensor cores, accumulators in tensor memory)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, kind::f16 (fp16/bf16 in, fp32 accumulate).
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once all previously issued MMAs have completed.
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// Same, arriving on the barrier at this offset in every CTA of `cta_mask`.
__device__ __forceinline__ void umma_commit_mcast(uint64_t *bar, uint16_t cta_mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(cta_mask)
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives lane
// (base_lane + i), columns [col, col + 32).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor for a K-major tile stored with the 128-byte
// swizzle TMA produces: rows of 128 B, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);  // start address, 16-byte units
  d |= static_cast<uint64_t>(1) << 16;                       // leading byte offset (unused with swizzle)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;               // stride byte offset: 8 rows * 128 B
  d |= static_cast<uint64_t>(1) << 46;                       // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;                       // layout: SWIZZLE_128B
  return d;
}
// Instruction descriptor, kind::f16: fp16 A/B (K-major), fp32 D, shape M x N.
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {
  return (1u << 4)                               // D format: F32
         | (0u << 7) | (0u << 10)                // A, B format: F16
         | (0u << 15) | (0u << 16)               // A, B major: K
         | (static_cast<uint32_t>(n >> 3) << 17) // N / 8
         | (static_cast<uint32_t>(m >> 4) << 24);  // M / 16
}

// ---------------------------------------------------------------------------
// Cross-GPU signalling from inside a kernel (see SignalArgs / WaitArgs in kernels.h).
// Both helpers contain a block-wide barrier: every thread of the block must call them.
// ---------------------------------------------------------------------------
template <class SignalT>
__device__ __forceinline__ void signal_peers_when_grid_done(const SignalT &sg) {
  if (sg.world <= 0) return;
  __threadfence_system();  // this thread's stores (peer memory included) precede the arrival below
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned blocks = gridDim.x * gridDim.y * gridDim.z;
    const unsigned prev = atomicAdd(sg.done_counter, 1u);
    if (prev == blocks - 1) {  // the last block of the launch
      *sg.done_counter = 0;    // ready for the next launch on this stream
      __threadfence_system();
      if (sg.mc_flags != nullptr)   // NVLS: the switch replicates the store into every rank's flag array
        asm volatile("multimem.st.release.sys.global.u32 [%0], %1;" ::"l"(sg.mc_flags + sg.rank), "r"(sg.value) : "memory");
      else
        for (int d = 0; d < sg.world; ++d)
          asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sg.peer_flags[d] + sg.rank), "r"(sg.value) : "memory");
    }
  }
}
template <class WaitT>
__device__ __forceinline__ void wait_for_peers(const WaitT &wa) {
  if (wa.world <= 0) return;
  if (static_cast<int>(threadIdx.x) < wa.world) {
    const long long t0 = clock64();
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(wa.flags + threadIdx.x) : "memory");
      if (static_cast<int>(v - wa.value) >= 0) break;
      if (clock64() - t0 > wa.timeout_cycles) {  // watchdog: a lost signal must not hang the GPU
        if (wa.status) atomicExch(wa.status, 600 + static_cast<int>(threadIdx.x));
        break;
      }
      __nanosleep(200);
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void named_barrier_sync(uint32_t id, uint32_t threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
#endif  // __CUDACC__

}  // namespace k
}  // namespace dmlp
