// CPU backends: the correctness oracle and the in-process multi-"rank" engine.
//
// The reference ecosystem shows five ways of distributing the same brute-force
// k-NN job: engine.cpp's 2-D process grid, bench_1's 1-D data
// shard with one batched gather, bench_2/3's 1-D data shard with a per-query
// tree reduction through a user-defined merge operator, bench_4's replicated
// dataset with a dynamic query farm, and bench.debug's serial kd-tree.  Each is
// available here as a `Strategy`, with "ranks" realised as host threads that
// only talk through explicit per-rank mailboxes, so the partition / padding /
// merge logic that the GPU paths reuse can be tested without MPI or GPUs.
#ifndef DMLP_CPU_H_
#define DMLP_CPU_H_

#include <string>

#include "dmlp/views.h"

namespace dmlp {

enum class Strategy {
  kSerial,           ///< one rank, brute force, threads over queries (oracle)
  kGrid2D,           ///< rows x cols grid (reference: engine.cpp:40-57)
  kShardGather,      ///< 1-D data shards, one gather, root merge (bench_1)
  kShardTreeReduce,  ///< 1-D data shards, pairwise merge-operator tree (bench_2/3)
  kQueryFarm,        ///< replicated data, dynamic query farm (bench_4)
  kKdTree,           ///< serial exact kd-tree (bench.debug)
};

/// Parse "serial|grid|gather|tree|farm|kdtree"; returns false on unknown names.
bool parse_strategy(const std::string &name, Strategy *out);
const char *strategy_name(Strategy s);

struct CpuOptions {
  Strategy strategy = Strategy::kSerial;
  int ranks = 1;    ///< number of emulated ranks (threads)
  int rows = 0;     ///< grid rows (data shards); 0 = derive from `ranks`
  int cols = 0;     ///< grid cols (query shards); 0 = derive from `ranks`
  int threads = 0;  ///< worker threads for kSerial; 0 = hardware concurrency
};

/// Squared Euclidean distance with the reference's exact rounding sequence:
/// fp64, strictly sequential `sum += (a[i]-b[i])*(a[i]-b[i])`, no FMA contraction
/// (reference: engine.cpp:12-18).  Every backend's final ordering is defined by
/// the bits this function returns.
double exact_sqdist(const double *a, const double *b, int dim);

/// Exact top-k of `data` for one query row, ascending in `closer` order and
/// padded to k with sentinels.  `out` must hold k entries.
void local_topk(const DatasetView &data, const double *query, int k, Neighbor *out);

/// The same for `nq` queries at once (rows `queries + i * data.a`, lists `outs[i]` of `k[i]`
/// entries): data rows are visited in groups of eight, transposed so that one SIMD register
/// holds the same attribute of eight points, and every group is used by all queries of a
/// block (64) before the next one is touched.  Each (query, point) pair still sees the exact
/// sequential sub / mul / add sequence of exact_sqdist, so the lists are bit-identical to
/// local_topk's; the dataset is streamed once per block instead of once per query.
void local_topk_many(const DatasetView &data, const double *queries, const int *k, int nq, Neighbor *const *outs);

/// The associative, commutative reduction operator of the tree strategy: merge
/// two `closer`-sorted k-lists and keep the first k (bench_2 @0xbc70 equivalent).
void merge_topk(const Neighbor *x, const Neighbor *y, int k, Neighbor *out);

/// Vote, checksum and append one query's final sorted list to `out`.
void finalize_query(const Neighbor *nb, int k, ResultSet *out);

/// Run the whole job on the CPU.  Results are indexed by query row.
void knn_cpu(const DatasetView &data, const QueryView &queries, const CpuOptions &opt,
             ResultSet *out);

}  // namespace dmlp

#endif  // DMLP_CPU_H_
