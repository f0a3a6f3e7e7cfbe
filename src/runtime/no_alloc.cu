// dmlp_knn_local_static: the local pipeline for callers that cannot allocate.
//
// Some hosts of this library have no way to give the pipeline a scratch buffer:
// a plug-in that runs inside another program's CUDA context under a "no
// allocation after start-up" rule, a real-time or safety-certified loop where
// cudaMalloc is banned because its latency and failure modes are unbounded, a
// process working inside a fixed memory reservation made by someone else.  For
// them the scratch lives inside the library image: one file-scope device array,
// reserved on every GPU that loads the library for as long as the process lives.
//
// The array's name in host code is only a handle for the runtime, not a device
// address, so each device's real address is looked up once with
// cudaGetSymbolAddress (with that device current) and remembered.  That lookup
// happens on the FIRST call made on a device; make that first call during
// start-up, before any stream capture begins on that device.
#include <mutex>

#include "dmlp_knn.h"

#include "../kernels/common.cuh"

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[DMLP_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_setup_lock;
void *g_scratch_addr[kMaxDevices] = {};

// Device address of the built-in scratch on the current device `dev`.
int resolve_scratch(int dev, void **out) {
  std::lock_guard<std::mutex> hold(g_setup_lock);
  if (g_scratch_addr[dev] == nullptr) {
    void *p = nullptr;
    DMLP_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
    DMLP_REQUIRE(p != nullptr, "knn_local_static: the runtime returned no address for the built-in scratch");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace

extern "C" int dmlp_knn_local_static(const double *q_attrs, const int *k, int q, const double *x_attrs,
                                     const int *x_labels, int n, int a, int id_base, int kmax, double *out_dist,
                                     int *out_id, int *out_label, int *flags, int *num_flagged, void *stream) {
  DMLP_REQUIRE(q >= 0 && n >= 0 && a >= 0 && kmax >= 1, "knn_local_static: negative size or kmax < 1");
  int dev = 0, sms = 0;
  DMLP_CUDA_OK(cudaGetDevice(&dev));
  DMLP_REQUIRE(dev >= 0 && dev < kMaxDevices, "knn_local_static: device index beyond the per-device address table");
  DMLP_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t need = dmlp_knn_local_workspace_bytes(q, n, a, sms, kmax);
  if (need > static_cast<size_t>(DMLP_STATIC_SCRATCH_BYTES)) {
    dmlp::k::set_error("knn_local_static: this problem needs " + std::to_string(need) +
                       " bytes of scratch but the built-in scratch holds " +
                       std::to_string(static_cast<size_t>(DMLP_STATIC_SCRATCH_BYTES)) +
                       "; use dmlp_knn_local with a caller-owned workspace");
    return -3;
  }
  void *scratch = nullptr;
  if (int rc = resolve_scratch(dev, &scratch)) return rc;
  // The scratch still holds whatever the previous call left in it.  Nothing in it
  // needs clearing here: the pipeline re-initialises the range / row-maximum
  // accumulators at the head of the scratch with its own reset kernels on every
  // call and zeroes the flagged-query counter (a caller output), and every other
  // region (packed rows, norms, candidate lists) is fully rewritten before it is
  // read.  Two calls that overlap on the GPU would both write it, so they race.
  return dmlp_knn_local(q_attrs, k, q, x_attrs, x_labels, n, a, id_base, kmax, /*range=*/nullptr, scratch,
                        static_cast<size_t>(DMLP_STATIC_SCRATCH_BYTES), out_dist, out_id, out_label, flags,
                        num_flagged, stream);
}
