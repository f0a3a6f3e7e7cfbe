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

}  // namespace

extern "C" {

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
