// Structure-of-arrays views and result containers shared by every backend.
//
// The reference moves array-of-structs `DataPoint`/`Query` objects and repacks
// them into flat id/label/attr arrays right before each MPI_Scatterv
// (reference: engine.cpp:79-106, 154-187).  Here the flat form is the native
// one: parsers and generators write it directly (into pinned host memory when a
// GPU backend is active), and the AoS types in common.h are only an API shim.
#ifndef DMLP_VIEWS_H_
#define DMLP_VIEWS_H_

#include <cstdint>
#include <vector>

namespace dmlp {

/// Sentinel used to pad a shard that holds fewer than k points.  Matches the
/// staff baselines: distance 1e18, id -1.
constexpr double kSentinelDist = 1e18;
constexpr int kSentinelId = -1;

/// Read-only view of `n` labelled points with `a` attributes each.
/// `attrs` is row-major `n x a`.  Global id of row r is `id_base + r`.
struct DatasetView {
  int64_t n = 0;
  int a = 0;
  const double *attrs = nullptr;
  const int *labels = nullptr;
  int id_base = 0;
  /// Optional: the same values as integers i with attrs[j] == double(i) / 1e6 exactly (the text
  /// grammar's six decimals, Problem::encode_decimal6).  A GPU backend then moves 4 bytes per
  /// value over PCIe instead of 8 and rebuilds the doubles bit for bit on the device.
  const int32_t *attrs_i32 = nullptr;
};

/// Read-only view of `q` queries.  Global id of row r is `id_base + r`.
struct QueryView {
  int64_t q = 0;
  int a = 0;
  const double *attrs = nullptr;
  const int *k = nullptr;
  int id_base = 0;
  const int32_t *attrs_i32 = nullptr;  ///< optional decimal6 form, see DatasetView
};

/// One candidate neighbour.  16 bytes; this is the "wire type" exchanged
/// between shards (reference: engine.cpp:214-228 registers the same triple as
/// an MPI struct datatype).
struct Neighbor {
  double dist;
  int id;
  int label;
};

/// Total order used both for selection and for reporting: ascending distance,
/// ties broken by the larger id.
inline bool closer(const Neighbor &x, const Neighbor &y) {
  return x.dist < y.dist || (x.dist == y.dist && x.id > y.id);
}

/// Per-query results in CSR form: neighbours of query i live in
/// [offset[i], offset[i+1]) of `dist`/`ids`, already in reporting order.
struct ResultSet {
  std::vector<int64_t> offset;
  std::vector<double> dist;
  std::vector<int> ids;
  std::vector<int> label;          ///< predicted label per query
  std::vector<uint64_t> checksum;  ///< release-mode checksum per query
  void clear() {
    offset.clear(); dist.clear(); ids.clear(); label.clear(); checksum.clear();
  }
};

/// 64-bit fold printed by reportResult (reference: common.cpp:59-68).  The basis
/// is the reference's literal, which is NOT the standard FNV-1a offset basis.
inline uint64_t result_checksum(int label, const int *ids, int64_t k) {
  uint64_t h = 1469598103934665603ULL;
  h ^= static_cast<uint64_t>(static_cast<long long>(label));
  h *= 1099511628211ULL;
  for (int64_t i = 0; i < k; ++i) {
    h ^= static_cast<uint64_t>(static_cast<long long>(ids[i] + 1));
    h *= 1099511628211ULL;
  }
  return h;
}

/// Majority vote over `k` neighbour labels; ties go to the larger label, an
/// empty neighbourhood yields -1 (reference: engine.cpp:320-332).  Sentinel
/// neighbours (id < 0) do not vote.
int vote_label(const Neighbor *nb, int64_t k);

/// Block partition of `total` items over `parts` owners with the remainder given
/// to owner 0 (reference: engine.cpp:62-63, 136-137).
struct BlockPartition {
  int64_t total = 0;
  int parts = 1;
  int64_t count(int p) const {
    return total / parts + (p == 0 ? total % parts : 0);
  }
  int64_t begin(int p) const {
    return p == 0 ? 0 : total % parts + p * (total / parts);
  }
};

/// Choose a near-square `rows x cols` factorisation with rows >= cols, the same
/// shape MPI_Dims_create hands the reference (reference: engine.cpp:40-44).
void choose_grid(int ranks, int *rows, int *cols);

}  // namespace dmlp

#endif  // DMLP_VIEWS_H_
