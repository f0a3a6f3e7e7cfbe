// NVSwitch multicast (NVLS) pages for the engine's cross-GPU flags.
//
// The K4 / K5 hand-shakes of a step end in "tell every peer": the last block of a launch stores
// an epoch into one word of every peer's flag array (common.cuh, signal_peers_when_grid_done).
// With a multicast object bound to one small page per GPU that is ONE store: the kernel writes
// `multimem.st.release.sys` to the multicast mapping and the switch replicates it into every
// GPU's page; waiters keep polling their own page through its ordinary (unicast) mapping.
//
// The reference has no counterpart (its signalling is MPI); SURVEY.md 5.9 names NVLS as the
// B200-native option.  Everything here is optional: when the driver, the GPUs or the fabric do
// not offer multicast, `McPages::create` says why and the engine keeps its per-peer stores.
#ifndef DMLP_ENGINE_MULTICAST_H_
#define DMLP_ENGINE_MULTICAST_H_

#include <cstddef>
#include <string>

namespace dmlp {
namespace gpu {

constexpr int kMcMaxDev = 8;

struct McPages {
  int ndev = 0;
  size_t bytes = 0;                      ///< per GPU (one multicast granule)
  unsigned char *local[kMcMaxDev] = {};  ///< GPU d's own page (unicast mapping; also readable by peers)
  unsigned char *mc = nullptr;           ///< multicast mapping: a store lands in every GPU's page
  bool ok = false;
  std::string why;                       ///< when !ok: what was missing

  /// One zeroed page on each of GPUs [0, ndev), bound to one multicast object.  Never throws.
  static McPages create(int ndev, size_t min_bytes);
  /// Unmap and release (the engine keeps its pages for the life of the process).
  void destroy();

 private:
  unsigned long long mc_handle_ = 0, mem_handle_[kMcMaxDev] = {};
};

}  // namespace gpu
}  // namespace dmlp

#endif  // DMLP_ENGINE_MULTICAST_H_
