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
// address, so the real address is looked up with cudaGetSymbolAddress, with the
// caller's device current.  A `__device__` variable exists once per CUDA CONTEXT
// (the primary context the runtime uses by default, or one a host application
// created itself and made current), so the looked-up address is remembered per
// context, not per device ordinal: a plug-in called from two contexts on the same
// GPU gets each context's own copy of the array.  The lookup happens on the FIRST
// call made in a context; make that first call during start-up, before any stream
// capture begins there.
#include <mutex>

#include "dmlp_knn.h"

#include "../kernels/common.cuh"

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[DMLP_STATIC_SCRATCH_BYTES];

typedef CUresult (*CtxGetCurrentFn)(CUcontext *);

constexpr int kMaxContexts = 64;
struct ContextSlot {
  CUcontext ctx;
  void *addr;
};
std::mutex g_setup_lock;
ContextSlot g_slots[kMaxContexts] = {};
int g_num_slots = 0;

// cuCtxGetCurrent through the runtime's driver entry-point query (no link-time dependency
// on libcuda); nullptr if the driver does not offer it.
CtxGetCurrentFn ctx_get_current_fn() {
  static CtxGetCurrentFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<CtxGetCurrentFn>(p);
    else
      cudaGetLastError();
  }
  return fn;
}

CUcontext current_context() {
  CUcontext ctx = nullptr;
  CtxGetCurrentFn fn = ctx_get_current_fn();
  if (fn != nullptr && fn(&ctx) != CUDA_SUCCESS) ctx = nullptr;
  return ctx;
}

// Device address of the built-in scratch in the context that is current on this thread.
int resolve_scratch(void **out) {
  std::lock_guard<std::mutex> hold(g_setup_lock);
  CUcontext ctx = current_context();
  if (ctx != nullptr)
    for (int i = 0; i < g_num_slots; ++i)
      if (g_slots[i].ctx == ctx) {
        *out = g_slots[i].addr;
        return 0;
      }
  // first call in this context (or no way to tell contexts apart: then look it up every time)
  void *p = nullptr;
  DMLP_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
  DMLP_REQUIRE(p != nullptr, "knn_local_static: the runtime returned no address for the built-in scratch");
  if (ctx == nullptr) ctx = current_context();  // the lookup itself may have bound the primary context
  if (ctx != nullptr && g_num_slots < kMaxContexts) {
    g_slots[g_num_slots].ctx = ctx;
    g_slots[g_num_slots].addr = p;
    ++g_num_slots;
  }
  *out = p;
  return 0;
}

}  // namespace

extern "C" int dmlp_knn_local_static(const double *q_attrs, const int *k, int q, const double *x_attrs,
                                     const int *x_labels, int n, int a, int id_base, int kmax, double *out_dist,
                                     int *out_id, int *out_label, int *flags, int *num_flagged, void *stream) {
  DMLP_REQUIRE(q >= 0 && n >= 0 && a >= 0 && kmax >= 1, "knn_local_static: negative size or kmax < 1");
  int dev = 0, sms = 0;
  DMLP_CUDA_OK(cudaGetDevice(&dev));
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
  if (int rc = resolve_scratch(&scratch)) return rc;
  // The scratch still holds whatever the previous call left in it.  Nothing in it
  // needs clearing here: the pipeline re-initialises the range / row-maximum
  // accumulators at the head of the scratch with its own reset kernels on every
  // call and zeroes the flagged-query counter (a caller output); with at least one
  // datapoint every other region (packed rows, norms, candidate lists) is fully
  // rewritten before it is read, and with none (n == 0) the pipeline never reads the
  // scratch at all and writes padding lists.  Two calls that overlap on the GPU would
  // both write it, so they race.
  return dmlp_knn_local(q_attrs, k, q, x_attrs, x_labels, n, a, id_base, kmax, /*range=*/nullptr, scratch,
                        static_cast<size_t>(DMLP_STATIC_SCRATCH_BYTES), out_dist, out_id, out_label, flags,
                        num_flagged, stream);
}
