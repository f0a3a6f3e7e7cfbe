// Stand-in for builds without the CUDA toolkit (make NO_CUDA=1): the GPU
// backends report zero devices and the engine runs its CPU strategies.
#include <stdexcept>

#include "dmlp/gpu.h"

namespace dmlp {
namespace gpu {
int device_count() { return 0; }
HostAllocator pinned_allocator() { return malloc_allocator(); }
void prepare(int64_t, int64_t, int, int, const GpuOptions &) {}
void knn_gpu(const DatasetView &, const QueryView &, const GpuOptions &, ResultSet *, std::string *) {
  throw std::runtime_error("built without CUDA");
}
}  // namespace gpu
}  // namespace dmlp
