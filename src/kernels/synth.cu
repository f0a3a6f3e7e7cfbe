// K8 synthetic_fill: the generator's distribution produced on the device
// (reference: generate_input.py:13-21 -- labels uniform, attributes uniform in
// [lo, hi] printed with six decimals, k uniform in [minK, min(maxK, N)]).
// Counter-based, so any shard of any array can be filled independently, and
// bit-identical to the host twin in include/dmlp/io.h: the arithmetic below
// uses explicit round-to-nearest intrinsics so no FMA contraction can change it.
#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

namespace {

__device__ __forceinline__ unsigned long long synth_hash_dev(unsigned long long seed, unsigned long long stream,
                                                             unsigned long long index) {
  unsigned long long z = seed * 0x9E3779B97F4A7C15ULL + stream * 0xD1B54A32D192ED03ULL + index;
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

__device__ __forceinline__ double synth_attr_dev(unsigned long long h, double lo, double hi) {
  const double u = __dmul_rn(static_cast<double>(h >> 11), 1.0 / 9007199254740992.0);
  const double x = __dadd_rn(lo, __dmul_rn(__dsub_rn(hi, lo), u));
  const double scaled = __dmul_rn(x, 1e6);
  const long long i = static_cast<long long>(scaled < 0 ? __dsub_rn(scaled, 0.5) : __dadd_rn(scaled, 0.5));
  return __ddiv_rn(static_cast<double>(i), 1e6);
}

__global__ void synth_attrs_kernel(double *dst, long long count, long long elem_base, double lo, double hi,
                                   unsigned long long seed, unsigned long long stream_id) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride)
    dst[i] = synth_attr_dev(synth_hash_dev(seed, stream_id, static_cast<unsigned long long>(elem_base + i)), lo, hi);
}

__global__ void synth_int_kernel(int *dst, long long count, long long base, int lo, int span,
                                 unsigned long long seed, unsigned long long stream_id) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride)
    dst[i] = lo + static_cast<int>(synth_hash_dev(seed, stream_id, static_cast<unsigned long long>(base + i)) %
                                   static_cast<unsigned long long>(span));
}

unsigned blocks_for(long long count) {
  long long b = (count + 255) / 256;
  return static_cast<unsigned>(b > 148 * 16 ? 148 * 16 : (b < 1 ? 1 : b));
}

}  // namespace

int launch_synth_attrs(double *dst, long long rows, int a, long long row_base, double lo, double hi,
                       unsigned long long seed, unsigned long long stream_id, cudaStream_t s) {
  const long long count = rows * a;
  if (count <= 0) return 0;
  synth_attrs_kernel<<<blocks_for(count), 256, 0, s>>>(dst, count, row_base * a, lo, hi, seed, stream_id);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_synth_labels(int *dst, long long rows, long long row_base, int num_labels, unsigned long long seed,
                        cudaStream_t s) {
  if (rows <= 0) return 0;
  DMLP_REQUIRE(num_labels > 0, "synth: num_labels must be positive");
  synth_int_kernel<<<blocks_for(rows), 256, 0, s>>>(dst, rows, row_base, 0, num_labels, seed, 1);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_synth_k(int *dst, long long rows, long long row_base, int min_k, int max_k, unsigned long long seed,
                   cudaStream_t s) {
  if (rows <= 0) return 0;
  DMLP_REQUIRE(max_k >= min_k && min_k >= 0, "synth: need 0 <= min_k <= max_k");
  synth_int_kernel<<<blocks_for(rows), 256, 0, s>>>(dst, rows, row_base, min_k, max_k - min_k + 1, seed, 3);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
