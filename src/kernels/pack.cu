// K7 pack_norms: fp64 rows -> centred, scaled fp16 rows + row norms + rounding-error norms.
//
// Replaces the reference's AoS->SoA repack before its scatters (reference:
// engine.cpp:79-96, 154-184).  The fp16 rows feed the tcgen05 filter GEMM; the
// norms turn its dot products into distances; the error norms make the filter
// certifiable (see rerank.cu).  A value v is mapped to x' = (v - c) * s with c
// the midpoint of the global value range and s a power of two chosen so that
// |x'| < 2^15 (fp16 keeps 11 significant bits at every magnitude up to 65504).
#include "common.cuh"
#include "kernels.h"

namespace dmlp {
namespace k {

__device__ __forceinline__ void atomic_max_double(double *addr, double v) {
  unsigned long long *a = reinterpret_cast<unsigned long long *>(addr);
  unsigned long long old = *a, assumed;
  do {
    assumed = old;
    if (__longlong_as_double(static_cast<long long>(assumed)) >= v) break;
    old = atomicCAS(a, assumed, static_cast<unsigned long long>(__double_as_longlong(v)));
  } while (assumed != old);
}

__global__ void stats_reset_kernel(PackStats *st) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    st->neg_lo = -INFINITY;
    st->hi = -INFINITY;
    st->max_norm = 0.f;
    st->max_err = 0.f;
  }
}

// Range of all values: max(-v) and max(v), reduced per block then one atomic.
__global__ void __launch_bounds__(256) stats_kernel(const double *__restrict__ src, long long count,
                                                    PackStats *st) {
  double hi = -INFINITY, nlo = -INFINITY;
  long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride) {
    double v = src[i];
    hi = fmax(hi, v);
    nlo = fmax(nlo, -v);
  }
  for (int o = 16; o > 0; o >>= 1) {
    hi = fmax(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
    nlo = fmax(nlo, __shfl_xor_sync(0xFFFFFFFFu, nlo, o));
  }
  __shared__ double s_hi[8], s_nlo[8];
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_hi[warp] = hi; s_nlo[warp] = nlo; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) { hi = fmax(hi, s_hi[w]); nlo = fmax(nlo, s_nlo[w]); }
    atomic_max_double(&st->hi, hi);
    atomic_max_double(&st->neg_lo, nlo);
  }
}

// One warp per row, grid-stride over rows.  Lanes own adjacent column pairs so
// the fp16 stores are 4-byte, 128 B per warp instruction.
template <bool I32>
__global__ void __launch_bounds__(256) pack_kernel(const void *__restrict__ src_v, long long rows, int a,
                                                   int a_pad, const PackStats *__restrict__ range,
                                                   __half *__restrict__ dst, float *__restrict__ norm2,
                                                   float *__restrict__ errn, PackStats *maxima, const WaitArgs wait,
                                                   const SignalArgs signal) {
  // (fused all-gathers) peers may still be reading the previous step's copy of these rows
  wait_for_peers(wait);
  const ScaleParams sp = derive_scale(range->neg_lo, range->hi);
  const int lane = threadIdx.x & 31;
  const long long warp_global = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long warps_total = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  // I32: the rows are resident in the decimal6 form (int32, value = i / 1e6), half the bytes to read
  const double *src = static_cast<const double *>(src_v);
  const int *src_i = static_cast<const int *>(src_v);
  const bool vec_ok = (a & 1) == 0 && (reinterpret_cast<uintptr_t>(src_v) & (I32 ? 7 : 15)) == 0;
  float my_max_norm = 0.f, my_max_err = 0.f;

  for (long long r = warp_global; r < rows; r += warps_total) {
    const double *in = src + r * a;
    const int *in_i = src_i + r * a;
    __half2 *out = reinterpret_cast<__half2 *>(dst + r * a_pad);
    double nsum = 0.0, esum = 0.0, xsum = 0.0;
    auto one = [&](double v0, double v1, bool in0, bool in1, int c) {
      double x0 = in0 ? __dmul_rn(__dsub_rn(v0, sp.center), sp.scale) : 0.0;
      double x1 = in1 ? __dmul_rn(__dsub_rn(v1, sp.center), sp.scale) : 0.0;
      __half h0 = __double2half(x0), h1 = __double2half(x1);
      double f0 = static_cast<double>(__half2float(h0)), f1 = static_cast<double>(__half2float(h1));
      // explicit operations: every instantiation of the kernel (fp64 or decimal6 source, block or tail
      // loop) must round the same way, or the same rows would get different norms
      double d0 = __dsub_rn(x0, f0), d1 = __dsub_rn(x1, f1);
      nsum = __fma_rn(f1, f1, __fma_rn(f0, f0, nsum));
      esum = __fma_rn(d1, d1, __fma_rn(d0, d0, esum));
      xsum = __fma_rn(x1, x1, __fma_rn(x0, x0, xsum));
      out[c >> 1] = __halves2half2(h0, h1);
    };
    int c0 = 0;
    if (vec_ok) {
      // whole 512-column blocks: eight independent 16-byte loads per lane are issued before any of
      // them is used (8 KB in flight per warp), each a fully coalesced 512-byte warp access
      for (; c0 + 512 <= a; c0 += 512) {
        if (I32) {
          int2 t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = __ldcs(reinterpret_cast<const int2 *>(in_i + c0 + j * 64 + lane * 2));
#pragma unroll
          for (int j = 0; j < 8; ++j) one(dec6_to_double(t[j].x), dec6_to_double(t[j].y), true, true, c0 + j * 64 + lane * 2);
        } else {
          double2 t[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) t[j] = __ldcs(reinterpret_cast<const double2 *>(in + c0 + j * 64 + lane * 2));
#pragma unroll
          for (int j = 0; j < 8; ++j) one(t[j].x, t[j].y, true, true, c0 + j * 64 + lane * 2);
        }
      }
    }
#pragma unroll 2
    for (int c = c0 + lane * 2; c < a_pad; c += 64) {
      double v0 = 0.0, v1 = 0.0;
      bool in0 = c < a, in1 = c + 1 < a;
      if (I32) {
        if (in0) v0 = dec6_to_double(in_i[c]);
        if (in1) v1 = dec6_to_double(in_i[c + 1]);
      } else if (vec_ok && in1) {
        double2 t = *reinterpret_cast<const double2 *>(in + c);
        v0 = t.x; v1 = t.y;
      } else {
        if (in0) v0 = in[c];
        if (in1) v1 = in[c + 1];
      }
      one(v0, v1, in0, in1, c);
    }
    for (int o = 16; o > 0; o >>= 1) {
      nsum += __shfl_xor_sync(0xFFFFFFFFu, nsum, o);
      esum += __shfl_xor_sync(0xFFFFFFFFu, esum, o);
      xsum += __shfl_xor_sync(0xFFFFFFFFu, xsum, o);
    }
    if (lane == 0) {
      // error norm: fp16 rounding plus the fp64 rounding of (v - c) * s, rounded up
      double e = (sqrt(esum) + 4.0 * 2.220446049250313e-16 * sqrt(xsum)) * (1.0 + 1e-9);
      float ef = __double2float_ru(e);
      float nf = __double2float_rn(nsum);
      norm2[r] = nf;
      errn[r] = ef;
      my_max_norm = fmaxf(my_max_norm, __double2float_ru(sqrt(nsum)));
      my_max_err = fmaxf(my_max_err, ef);
    }
  }
  if (lane == 0 && maxima != nullptr) {
    // non-negative floats order like their bit patterns
    atomicMax(reinterpret_cast<int *>(&maxima->max_norm), __float_as_int(my_max_norm));
    atomicMax(reinterpret_cast<int *>(&maxima->max_err), __float_as_int(my_max_err));
  }
  signal_peers_when_grid_done(signal);  // "my packed rows are in place"
}

// Wire format "decimal6": the text grammar carries six decimals, so a value whose
// scaled integer fits 32 bits travels as that integer (half the PCIe bytes of an
// fp64) and is expanded here.  (double)i / 1e6 with IEEE division is correctly
// rounded, i.e. exactly the double a text parser produces for that decimal.
__global__ void __launch_bounds__(256) expand_decimal6_kernel(const int *__restrict__ src, double *__restrict__ dst,
                                                              long long count) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride)
    dst[i] = dec6_to_double(src[i]);   // == __ddiv_rn((double)src[i], 1e6), see common.cuh
}

// Checks that every value survives the decimal6 round trip and emits the integers.
__global__ void __launch_bounds__(256) encode_decimal6_kernel(const double *__restrict__ src, int *__restrict__ dst,
                                                              long long count, int *bad) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  int local_bad = 0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride) {
    const double v = src[i];
    const double scaled = rint(__dmul_rn(v, 1e6));
    int iv = 0;
    if (fabs(scaled) < 2147483647.0) {
      iv = static_cast<int>(scaled);
      if (__ddiv_rn(static_cast<double>(iv), 1e6) != v) local_bad = 1;
    } else {
      local_bad = 1;
    }
    dst[i] = iv;
  }
  if (local_bad) atomicExch(bad, 1);
}

// Bit-packed decimal6 wire: a value i with |i| < 2^(bits-1) travels as the `bits`-bit unsigned
// number i + 2^(bits-1); a row of `a` values is a little-endian bit stream padded to whole 32-bit
// words (`words` per row).  Values in [-100, 100] with six decimals need 28 bits: 12.5 % fewer
// bytes over PCIe than int32, 56 % fewer than fp64.
__global__ void __launch_bounds__(256) pack_bits_kernel(const int *__restrict__ src, long long rows, int a, int bits,
                                                        int words, unsigned *__restrict__ dst) {
  const long long total = rows * words;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const unsigned long long mask = (1ULL << bits) - 1ULL;
  const unsigned bias = 1u << (bits - 1);
  for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
    const long long r = t / words;
    const int w = static_cast<int>(t - r * words);
    const int *row = src + r * a;
    const long long bit0 = 32LL * w;
    unsigned word = 0;
    for (int c = static_cast<int>(bit0 / bits); c < a && static_cast<long long>(c) * bits < bit0 + 32; ++c) {
      const unsigned long long u = static_cast<unsigned long long>(static_cast<unsigned>(row[c]) + bias) & mask;
      const long long sh = static_cast<long long>(c) * bits - bit0;   // position of the value's bit 0 in this word
      word |= sh >= 0 ? static_cast<unsigned>(u << sh) : static_cast<unsigned>(u >> (-sh));
    }
    dst[t] = word;
  }
}

template <bool TO_F64>
__global__ void __launch_bounds__(256) unpack_bits_kernel(const unsigned *__restrict__ src, long long rows, int a,
                                                          int bits, int words, int rows_per_chunk,
                                                          int *__restrict__ dst_i, double *__restrict__ dst_d) {
  const unsigned long long mask = (1ULL << bits) - 1ULL;
  const unsigned bias = 1u << (bits - 1);
  const long long chunks = (rows + rows_per_chunk - 1) / rows_per_chunk;
  for (long long ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
    const long long r0 = ch * rows_per_chunk;
    const unsigned nrows = static_cast<unsigned>(rows - r0 < rows_per_chunk ? rows - r0 : rows_per_chunk);
    const unsigned *in = src + r0 * words;
    const long long out0 = r0 * a;
    for (unsigned idx = threadIdx.x; idx < nrows * static_cast<unsigned>(a); idx += 256) {   // 32-bit index math
      const unsigned r = idx / static_cast<unsigned>(a), c = idx - r * static_cast<unsigned>(a);
      const unsigned off = c * static_cast<unsigned>(bits);
      const unsigned w = off >> 5, sh = off & 31u;
      const unsigned *row = in + r * static_cast<unsigned>(words);
      unsigned long long v = row[w];
      if (sh + bits > 32) v |= static_cast<unsigned long long>(row[w + 1]) << 32;   // (w + 1 < words whenever bits are left)
      const int i = static_cast<int>(static_cast<unsigned>((v >> sh) & mask) - bias);
      if (TO_F64) dst_d[out0 + idx] = dec6_to_double(i);
      else dst_i[out0 + idx] = i;
    }
  }
}

int launch_pack_bits(const int *src, long long rows, int a, int bits, int words, unsigned *dst, cudaStream_t s) {
  if (rows <= 0) return 0;
  DMLP_REQUIRE(bits >= 2 && bits <= 32 && static_cast<long long>(words) * 32 >= static_cast<long long>(a) * bits,
               "pack_bits: bits must be in [2, 32] and the row must fit its words");
  long long blocks = (rows * words + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pack_bits_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, rows, a, bits, words, dst);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_unpack_bits(const unsigned *src, long long rows, int a, int bits, int words, int *dst_i32, double *dst_f64,
                       cudaStream_t s) {
  if (rows <= 0) return 0;
  DMLP_REQUIRE(bits >= 2 && bits <= 32 && static_cast<long long>(words) * 32 >= static_cast<long long>(a) * bits,
               "unpack_bits: bits must be in [2, 32] and the row must fit its words");
  DMLP_REQUIRE((dst_i32 != nullptr) != (dst_f64 != nullptr), "unpack_bits: exactly one destination");
  DMLP_REQUIRE(a >= 1 && a <= (1 << 20), "unpack_bits: 1 <= a <= 2^20");
  const int rpc = a >= 2048 ? 1 : 2048 / a;          // about 2048 values (8 per thread) per block and pass
  long long blocks = (rows + rpc - 1) / rpc;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (dst_f64 != nullptr)
    unpack_bits_kernel<true><<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, rows, a, bits, words, rpc, nullptr,
                                                                           dst_f64);
  else
    unpack_bits_kernel<false><<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, rows, a, bits, words, rpc, dst_i32,
                                                                            nullptr);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_expand_decimal6(const int *src, double *dst, long long count, cudaStream_t s) {
  if (count <= 0) return 0;
  long long blocks = (count + 256 * 8 - 1) / (256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  expand_decimal6_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, dst, count);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_encode_decimal6(const double *src, int *dst, long long count, int *bad, cudaStream_t s) {
  if (count <= 0) return 0;
  long long blocks = (count + 256 * 8 - 1) / (256 * 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  encode_decimal6_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, dst, count, bad);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

__global__ void __launch_bounds__(256) maxima_from_norms_kernel(const float *__restrict__ norm2,
                                                                const float *__restrict__ errn, long long rows,
                                                                PackStats *maxima) {
  float mn = 0.f, me = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < rows; i += stride) {
    mn = fmaxf(mn, norm2[i]);
    me = fmaxf(me, errn[i]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmaxf(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
    me = fmaxf(me, __shfl_xor_sync(0xFFFFFFFFu, me, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(reinterpret_cast<int *>(&maxima->max_norm), __float_as_int(__fsqrt_ru(mn) * (1.0f + 1e-6f)));
    atomicMax(reinterpret_cast<int *>(&maxima->max_err), __float_as_int(me));
  }
}

int launch_maxima_from_norms(const float *norm2, const float *errn, long long rows, PackStats *maxima,
                             cudaStream_t s) {
  if (rows <= 0) return 0;
  long long blocks = (rows + 256 * 4 - 1) / (256 * 4);
  if (blocks > 148 * 4) blocks = 148 * 4;
  maxima_from_norms_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(norm2, errn, rows, maxima);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_stats_reset(PackStats *st, cudaStream_t s) {
  stats_reset_kernel<<<1, 32, 0, s>>>(st);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_stats(const double *src, long long count, PackStats *st, cudaStream_t s) {
  if (count <= 0) return 0;
  long long blocks = (count + 256 * 8 - 1) / (256 * 8);
  if (blocks > 148 * 8) blocks = 148 * 8;
  stats_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, count, st);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_pack(const void *src, long long rows, int a, int a_pad, const PackStats *range, __half *dst,
                float *norm2, float *errn, PackStats *maxima, cudaStream_t s, const WaitArgs *wait,
                const SignalArgs *signal, bool src_i32) {
  const WaitArgs wa = wait ? *wait : WaitArgs();
  const SignalArgs sg = signal ? *signal : SignalArgs();
  if (rows <= 0 && sg.world <= 0) return 0;
  if (rows < 0) rows = 0;
  DMLP_REQUIRE(a_pad >= a && (a_pad % 8) == 0, "pack: a_pad must be a multiple of 8 and >= a");
  long long blocks = (rows + 7) / 8;  // 8 warps per block
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  if (src_i32)
    pack_kernel<true><<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, rows, a, a_pad, range, dst, norm2, errn, maxima,
                                                                   wa, sg);
  else
    pack_kernel<false><<<static_cast<unsigned>(blocks), 256, 0, s>>>(src, rows, a, a_pad, range, dst, norm2, errn,
                                                                    maxima, wa, sg);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace k
}  // namespace dmlp
