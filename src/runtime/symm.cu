// Symmetric peer memory and device-side signalling between GPUs.
//
// Replaces the reference's communication layer (MPI_Bcast / Scatterv / Gather /
// Barrier on host memory, reference: engine.cpp:33-35, 73-131, 147-209, 282-284):
// every rank allocates the same buffers, exports them with CUDA IPC, maps its
// peers' copies over NVLink, and the kernels themselves load and store through
// those mappings.  Flags are plain words in the same buffers: a producer
// finishes its data, fences at system scope and stores an epoch number into the
// consumer's flag word; the consumer spins on its own local memory.
#include "dmlp_knn.h"

#include <cstring>

#include "../kernels/common.cuh"

namespace {

__global__ void signal_kernel(uint32_t *const *peer_flags, int world, int my_rank, uint32_t value) {
  // one thread per destination rank
  const int d = threadIdx.x;
  if (d >= world) return;
  __threadfence_system();  // everything this GPU wrote before (earlier kernels included) is visible first
  uint32_t *dst = peer_flags[d] + my_rank;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(value) : "memory");
}

__global__ void wait_kernel(const uint32_t *flags, int world, uint32_t value, long long timeout_cycles,
                            int *status) {
  const int s = threadIdx.x;
  if (s >= world) return;
  const long long t0 = clock64();
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + s) : "memory");
    if (static_cast<int32_t>(v - value) >= 0) break;
    if (clock64() - t0 > timeout_cycles) {  // watchdog: a lost signal must not hang the GPU
      if (status) atomicExch(status, 1 + s);
      break;
    }
    __nanosleep(200);
  }
}

struct PeerGather {
  int world, rank;
  uint32_t epoch;
  unsigned char *peer_base[8];
  unsigned long long off_flags;
  unsigned long long part_off[8], part_len[8];
  int *status;
  long long timeout_cycles;
};

template <class T>
__device__ __forceinline__ void copy_units(unsigned char *dst_b, const unsigned char *src_b, unsigned long long n) {
  T *dst = reinterpret_cast<T *>(dst_b);
  const T *src = reinterpret_cast<const T *>(src_b);
  const long long count = static_cast<long long>(n);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < count; i += 4 * stride) {   // four independent loads in flight per thread
    T v0 = src[i], v1 = src[i + stride], v2 = src[i + 2 * stride], v3 = src[i + 3 * stride];
    dst[i] = v0; dst[i + stride] = v1; dst[i + 2 * stride] = v2; dst[i + 3 * stride] = v3;
  }
  for (; i < count; i += stride) dst[i] = src[i];
}

// All-gather of per-rank parts of one symmetric region: every GPU pulls the other
// ranks' parts over NVLink into the same offsets of its own buffer.  Peers are
// visited in a rank-rotated order so that no single GPU is everybody's first source.
__global__ void __launch_bounds__(256) peer_gather_kernel(const PeerGather g) {
  __shared__ int s_dead;
  if (threadIdx.x == 0) s_dead = 0;
  for (int j = 1; j < g.world; ++j) {
    const int p = (g.rank + j) % g.world;
    if (threadIdx.x == 0) {
      const uint32_t *flag = reinterpret_cast<const uint32_t *>(g.peer_base[g.rank] + g.off_flags) + p;
      const long long t0 = clock64();
      for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (static_cast<int32_t>(v - g.epoch) >= 0) break;
        if (clock64() - t0 > g.timeout_cycles) {
          if (g.status) atomicExch(g.status, 300 + p);
          s_dead = 1;
          break;
        }
        __nanosleep(100);
      }
    }
    __syncthreads();
    if (s_dead) return;
    const unsigned long long off = g.part_off[p], len = g.part_len[p];
    if (((off | len) & 15) == 0) copy_units<int4>(g.peer_base[g.rank] + off, g.peer_base[p] + off, len / 16);
    else if (((off | len) & 7) == 0) copy_units<long long>(g.peer_base[g.rank] + off, g.peer_base[p] + off, len / 8);
    else if (((off | len) & 3) == 0) copy_units<int>(g.peer_base[g.rank] + off, g.peer_base[p] + off, len / 4);
    else copy_units<unsigned char>(g.peer_base[g.rank] + off, g.peer_base[p] + off, len);
  }
}

}  // namespace

extern "C" {

int dmlp_peer_gather(const dmlp_peer_gather_args *a, void *stream) {
  DMLP_REQUIRE(a != nullptr && a->world >= 1 && a->world <= 8, "peer_gather: world must be in [1, 8]");
  if (a->world == 1) return 0;
  PeerGather g;
  g.world = a->world; g.rank = a->rank; g.epoch = a->epoch;
  for (int i = 0; i < a->world; ++i) {
    g.peer_base[i] = static_cast<unsigned char *>(a->peer_base[i]);
    g.part_off[i] = a->part_off[i];
    g.part_len[i] = a->part_len[i];
  }
  g.off_flags = a->off_flags;
  g.status = a->status;
  g.timeout_cycles = static_cast<long long>((a->timeout_seconds > 0 ? a->timeout_seconds : 20.0) * 1.5e9);
  int dev = 0, sms = 148;
  DMLP_CUDA_OK(cudaGetDevice(&dev));
  DMLP_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  peer_gather_kernel<<<sms * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(g);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int dmlp_symm_alloc(size_t bytes, void **ptr, unsigned char handle_out[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void *p = nullptr;
  DMLP_CUDA_OK(cudaMalloc(&p, bytes ? bytes : 256));
  DMLP_CUDA_OK(cudaMemset(p, 0, bytes ? bytes : 256));
  cudaIpcMemHandle_t h;
  DMLP_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle_out, &h, 64);
  *ptr = p;
  return 0;
}

int dmlp_symm_open(const unsigned char handle[64], void **ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  DMLP_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int dmlp_symm_close(void *ptr) {
  DMLP_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int dmlp_symm_free(void *ptr) {
  DMLP_CUDA_OK(cudaFree(ptr));
  return 0;
}

int dmlp_symm_signal(void *peer_flag_ptrs, int world, int my_rank, uint32_t value, void *stream) {
  DMLP_REQUIRE(world >= 1 && world <= 32, "symm_signal: world must be in [1, 32]");
  signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint32_t *const *>(peer_flag_ptrs), world,
                                                                 my_rank, value);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int dmlp_symm_wait(const void *local_flags, int world, uint32_t value, double timeout_seconds, int *status,
                   void *stream) {
  DMLP_REQUIRE(world >= 1 && world <= 32, "symm_wait: world must be in [1, 32]");
  const long long cycles = static_cast<long long>(timeout_seconds * 1.5e9);
  wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const uint32_t *>(local_flags), world, value,
                                                               cycles, status);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
