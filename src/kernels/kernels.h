// Internal launch interface of the sm_100a kernels (host side).  The public C
// ABI in include/dmlp_knn.h is a thin layer over these.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>

namespace dmlp {
namespace k {

/// Device-resident range and row maxima produced by the stats / pack kernels.
/// `neg_lo` and `hi` are both max-reductions so one MAX all-reduce merges ranks.
struct PackStats {
  double neg_lo;   ///< max over all values of -v
  double hi;       ///< max over all values of v
  float max_norm;  ///< max over packed rows of ||x^||
  float max_err;   ///< max over packed rows of ||x' - x^||
};

struct ScaleParams {
  double center;
  double scale;  ///< power of two
};

/// Midpoint centring and power-of-two scaling so that |(v - c) * s| < 2^15.
__host__ __device__ inline ScaleParams derive_scale(double neg_lo, double hi) {
  ScaleParams p;
  p.center = 0.0;
  p.scale = 1.0;
  double lo = -neg_lo;
  if (!(lo <= hi)) return p;  // empty or NaN range
  p.center = 0.5 * lo + 0.5 * hi;
  double m = fmax(hi - p.center, p.center - lo);
  if (!(m > 0.0) || !isfinite(m)) return p;
  int e = ilogb(m);  // m in [2^e, 2^(e+1))
  int sh = 14 - e;
  if (sh > 1000) sh = 1000;
  if (sh < -1000) sh = -1000;
  p.scale = scalbn(1.0, sh);
  return p;
}

/// Work decomposition of the filter GEMM (host decides, kernel follows).
struct TopkPlan {
  int q_rows;      ///< queries
  int n;           ///< local datapoints
  int a_pad;       ///< fp16 row length (multiple of 8)
  int kp;          ///< candidates kept per (query, chunk)
  int m_blocks;    ///< ceil(q_rows / 128)
  int n_tiles;     ///< ceil(n / 256)
  int group_m;     ///< row blocks that share one wave (= group_mc * cluster)
  int chunks;      ///< column chunks per query (lists per query)
  int k_blocks;    ///< ceil(a_pad / 64)
  int grid;        ///< persistent CTAs (= clusters * cluster)
  int pair;        ///< 1: CTA-pair MMA (cta_group::2, cluster of 2); 0: one CTA per tile
  int cluster;     ///< CTAs per cluster; they sweep the same data tiles via TMA multicast
  int mc_blocks;   ///< ceil(m_blocks / cluster): row-block groups handled by one cluster
  int group_mc;    ///< cluster rows per wave
  int sync;        ///< > 0: clusters sweeping the same chunk re-align every `sync` data tiles (keeps their tile reads within L2 reach)
  unsigned *sync_cnt;  ///< device counters for `sync`, one per (group, chunk), zeroed per launch
  int cand_chunks; ///< lists per query in the candidate buffer (>= chunks; several launches can share one buffer)
  int chunk_base;  ///< first list index this launch writes
  int row_base;    ///< added to the row ids this launch emits (the operand is a slice of a larger shard)
  int wide;        ///< pair kernel only: 1 = one query tile meets two data tiles per k-block (256 x 512 per pair; query tiles are read half as often)
  const unsigned *gate_cnt;  ///< set when the data rows are gathered from peers by the kernel itself: per-tile arrival counters the epilogue checks before it stages a tile's norms
  unsigned gate_epoch;
  int gap;         ///< wide only: >= 0 = skewed schedule, steps an accumulator rests between tiles; -1 = lock-step wide tile
  int halves;      ///< 2 (short rows, pair kernel): two epilogue warps per TMEM lane quarter, each scanning half of a tile's columns into a list of its own -- every (query, chunk) then yields two lists
  int lists;       ///< lists per query this plan writes (chunks * halves, times parts): what the candidate buffer and the re-rank are sized for
  int parts;       ///< > 1 (large k): the launch is split into `parts` launches over row ranges of the data, each writing `part_chunks` of the `chunks` lists -- a list keeps 32 keys, so k up to 1024 needs up to 128 lists per query
  int part_chunks; ///< lists per part (= lists when parts == 1)
  int part_kmax;   ///< the kmax each part is planned for
  int num_sms;     ///< what the plan was made for (the parts are planned at launch)
  int dyn;         ///< clusters take items from an atomic counter instead of the static round-robin: 1 always (switches the re-alignment off), 0 never, -1 (default) where the launch has no re-alignment anyway; single-GPU launches only (DMLP_TOPK_DYN)
  unsigned *dyn_counter;   ///< next item to hand out (zeroed per launch)
  unsigned *dyn_slots;     ///< [clusters][dyn_cap]: item + 2 of a cluster's i-th grab, 1 = no more items, 0 = not yet known
  int dyn_cap;
  unsigned long long *prof;  ///< diagnostics (DMLP_TOPK_PROF=1): per CTA 16 cycle counters, see dist_topk_prof_read
  int rot;         ///< rotation of the cluster-row order (multi-GPU: start with the local slice)
  int hint_q, hint_x;  ///< L2 eviction hints for query / data tiles: 0 normal, 1 first, 2 last
};
/// `kmax`: largest k any query asks for (0 = unknown); larger k gets more chunks, because
/// every chunk contributes at most 32 candidates per query.
TopkPlan make_topk_plan(int q_rows, int n, int a_pad, int kp, int num_sms, int kmax = 0);

constexpr int kMaxPeers = 8;
/// "When this launch has finished, tell the peers" (K5 signalling from inside the kernels):
/// every block bumps `done_counter` as it finishes; the last one stores `value` into word
/// `rank` of every peer's flag array with release semantics at system scope and resets the
/// counter.  world == 0: no signal.
struct SignalArgs {
  unsigned *peer_flags[kMaxPeers] = {};
  int world = 0, rank = 0;
  unsigned value = 0;
  unsigned *done_counter = nullptr;
  unsigned *mc_flags = nullptr;  ///< multicast mapping of the flag arrays (NVLS): one multimem store reaches every rank
};
/// "Do not start before the peers have told us": every block spins (acquire, system scope) until
/// each of the `world` words of the local flag array has reached `value`; a wait that exceeds
/// timeout_cycles writes 600 + the late rank to *status and carries on.  world == 0: no wait.
struct WaitArgs {
  const unsigned *flags = nullptr;
  int world = 0;
  unsigned value = 0;
  long long timeout_cycles = 0;
  int *status = nullptr;
};

// pack.cu
int launch_stats_reset(PackStats *st, cudaStream_t s);
int launch_stats(const double *src, long long count, PackStats *st, cudaStream_t s);
/// src: fp64 rows, or (src_i32) the same rows in decimal6 form: int32 i with value i / 1e6
int launch_pack(const void *src, long long rows, int a, int a_pad, const PackStats *range, __half *dst,
                float *norm2, float *errn, PackStats *maxima, cudaStream_t s, const WaitArgs *wait = nullptr,
                const SignalArgs *signal = nullptr, bool src_i32 = false);

/// Cross-GPU operand gather fused into the filter GEMM (K4).  Every rank owns
/// `slice_rows` query rows and has packed them into its own symmetric buffer;
/// a communication warp in each GEMM CTA pulls the other ranks' packed rows over
/// NVLink into the local copy just ahead of the tiles that consume them, then
/// pulls the fp64 rows, norms and k for the re-rank kernel that follows.
struct GatherArgs {
  int world = 1, rank = 0;
  int slice_rows = 0;          ///< padded rows per owner (multiple of 256)
  int a = 0;                   ///< fp64 row length
  unsigned epoch = 0;
  unsigned char *peer_base[kMaxPeers] = {};  ///< symmetric buffers as mapped here (own included)
  unsigned long long off_qh = 0, off_q64 = 0, off_qn = 0, off_qe = 0, off_k = 0, off_pack_flags = 0;
  unsigned *rb_count = nullptr;  ///< local, one counter per 128-row block, monotonically increasing
  int *status = nullptr;
  long long timeout_cycles = 0;
};

/// The same idea for the data operand (the reference's row Bcast of a data shard,
/// engine.cpp:129-131, fused into the GEMM): the ranks of a grid row each packed one part
/// of the shard; the communication warps pull the other parts' packed rows and norms tile
/// by tile in the order the sweep needs them (per-tile counters gate the TMA producers),
/// then the fp64 rows and labels for the re-rank.
struct DataGatherArgs {
  int world = 1, rank = 0;
  int a = 0;                   ///< fp64 row length
  unsigned epoch = 0;
  unsigned char *peer_base[kMaxPeers] = {};
  unsigned long long off_xh = 0, off_xn = 0, off_xe = 0, off_x64 = 0, off_labels = 0, off_pack_flags = 0;
  int part_begin[kMaxPeers + 1] = {};  ///< rank p packed rows [part_begin[p], part_begin[p + 1])
  unsigned *tile_count = nullptr;      ///< local, one counter per 256-row data tile, monotonically increasing
  int *status = nullptr;
  long long timeout_cycles = 0;
  SignalArgs done;                     ///< optional: "I have finished reading your parts", sent when the launch ends
  int x_units = 0;                     ///< 8-byte units per row of the region at off_x64 (a for fp64 rows, a / 2 for decimal6 rows; 0 = a)
};

/// Where the exact k-lists go (K5): straight into the owner rank's receive
/// buffer [world][slice_rows][kmax], through the peer mapping.
struct ScatterArgs {
  int world = 1, rank = 0;
  int slice_rows = 0;
  unsigned char *peer_base[kMaxPeers] = {};
  unsigned long long off_dist = 0, off_id = 0, off_label = 0;
};

int launch_expand_decimal6(const int *src, double *dst, long long count, cudaStream_t s);
/// bit-packed decimal6 wire format (pack.cu): `bits` per value, `words` 32-bit words per row
int launch_pack_bits(const int *src, long long rows, int a, int bits, int words, unsigned *dst, cudaStream_t s);
int launch_unpack_bits(const unsigned *src, long long rows, int a, int bits, int words, int *dst_i32, double *dst_f64,
                       cudaStream_t s);
int launch_encode_decimal6(const double *src, int *dst, long long count, int *bad, cudaStream_t s);

/// max row norm / error norm over packed rows, from the norm arrays (after a gather)
int launch_maxima_from_norms(const float *norm2, const float *errn, long long rows, PackStats *maxima, cudaStream_t s);

// dist_topk.cu  (cand: [q_rows][chunks][kp] keys, ascending per list)
int launch_dist_topk(const __half *qh, const __half *xh, const float *xnorm2, const TopkPlan &plan,
                     unsigned long long *cand, const GatherArgs *gather, cudaStream_t s,
                     const DataGatherArgs *data_gather = nullptr);

/// The work items of a plan in the order the kernel walks them (host replay of the kernel's
/// enumeration, for tests): 8 ints per item and CTA {cluster, cta rank, row block, chunk,
/// first tile, end tile, wave group, cluster rows in the group}.  Returns the item count
/// (which may exceed max_items; only max_items are stored).
int enumerate_topk_items(const TopkPlan &plan, int *out, int max_items);

/// Diagnostics of the last dist_topk launch made with DMLP_TOPK_PROF=1 (pair kernel): per CTA
/// {mma total, mma wait for operands, mma wait for a drained accumulator, producer total,
///  producer wait for a free stage, producer re-alignment wait, epilogue wait for a full
///  accumulator, epilogue total, epilogue norm staging, accumulator loads, scoring and appends,
///  list compaction, list flush, 3 unused} in SM cycles (16 values per CTA).  Returns the number of CTAs written.
int dist_topk_prof_read(unsigned long long *host_out, int max_ctas);
/// Number of re-alignment waits that timed out on the current device since the library was
/// loaded (synchronises the device).  0 in a healthy run.
int dist_topk_sync_timeouts(unsigned *out);

// rerank.cu
struct RerankArgs {
  const unsigned long long *cand;  // [q][chunks][kp]
  int chunks, kp;
  const double *q_attrs;           // [q][a] original fp64
  const double *x_attrs;           // [n][a]
  const int *x_attrs_i32 = nullptr;  // optional [n][a]: the same rows in decimal6 form (int32 i, value i / 1e6); read instead of x_attrs
  const int *x_labels;             // [n]
  const int *k;                    // [q]
  const float *q_norm2;            // [q]  ||q^||^2 (scaled units)
  const float *q_errn;             // [q]  ||dq||
  const PackStats *range;          // scale parameters
  const PackStats *x_maxima;       // max ||x^||, max ||dx|| over the local shard
  int q, n, a, a_pad;
  int id_base;                     // global id of local row 0
  int kmax;                        // output stride
  double *out_dist;                // [q][kmax]
  int *out_id;                     // [q][kmax] global ids, -1 padded
  int *out_label;                  // [q][kmax]
  int *flags;                      // [q] 0 = certified, else reason bits
  int *num_flagged;                // device counter
  ScatterArgs scatter;             // world > 1: lists go to the owners' receive buffers
  int *w_out = nullptr;            // optional [q]: survivors recomputed per query (diagnostics)
  int *final_label = nullptr;      // optional [q] (single data shard only): predicted label, and
  unsigned long long *final_checksum = nullptr;  //   [q] checksum -- the vote + checksum tail folded into the emit kernel
  int *flagged_rows = nullptr;     // optional [flagged_cap]: indices of flagged queries, slot = their order in num_flagged
  int flagged_cap = 0;
  SignalArgs signal;               // optional: the emit kernel tells the peers when the whole launch is done
  void *workspace = nullptr;       // >= rerank_workspace_bytes(q), 256-byte aligned
  size_t workspace_bytes = 0;
};
size_t rerank_workspace_bytes(int q, int kmax);
int launch_rerank(const RerankArgs &args, cudaStream_t s);
void set_rerank_debug_w(int *w_out);  // diagnostics: next launches record survivors per query here

// finalize.cu: merge `sources` sorted lists per query, vote, checksum
struct FinalizeArgs {
  const double *dist;   // [sources][q][kmax]
  const int *id;        // [sources][q][kmax]
  const int *label;     // [sources][q][kmax]
  long long source_stride;  // elements between sources (>= q*kmax)
  int sources, q, kmax;
  const int *k;         // [q]
  int *out_label;       // [q]
  unsigned long long *out_checksum;  // [q]
  int *out_id;          // [q][kmax] optional (nullptr to skip)
  double *out_dist;     // [q][kmax] optional
  void *workspace = nullptr;   // >= finalize_workspace_bytes(); only needed when sources > 1 and kmax > 2048
  size_t workspace_bytes = 0;
  const int *query_rows = nullptr;  // optional: only these rows are finalized (`count` entries at most,
  int count = 0;                    //   the first min(count, *count_dev) when count_dev is set: a count
  const int *count_dev = nullptr;   //   that only the device knows, e.g. the flagged queries of this step)
  WaitArgs wait;                    // optional: the lists come from peers; wait for their signals first
};
size_t finalize_workspace_bytes(int sources, int q, int kmax);
int launch_finalize(const FinalizeArgs &args, cudaStream_t s);

// exact.cu: brute-force fp64 path (any k), used for tiny problems, large k and
// queries the filter could not certify
struct ExactArgs {
  const double *q_attrs;  // [q][a]
  const double *x_attrs;  // [n][a]
  const int *x_attrs_i32 = nullptr;  // optional: the rows in decimal6 form, read instead of x_attrs
  const int *x_labels;
  const int *k;           // [q]
  const int *query_rows;  // optional [count] subset of query rows (nullptr = all q rows)
  int count;              // number of queries to process (an upper bound when count_dev is set)
  const int *count_dev = nullptr;  // optional device int: process min(count, *count_dev) queries -- lets a
                                   // repair pass be enqueued before anyone knows how many queries need it
  ScatterArgs scatter;    // world > 1: lists go to the owners' receive buffers (like RerankArgs)
  SignalArgs signal;      // optional: the merge kernel tells the peers when the whole launch is done
  int n, a, id_base, kmax;
  double *out_dist;       // [q][kmax] (written at the query's own row)
  int *out_id;
  int *out_label;
  void *workspace;        // 256-byte aligned; exact_workspace_bytes(n, count, kmax) holds every query at once,
  size_t workspace_bytes; // anything down to exact_workspace_bytes(n, 1, kmax) works (smaller batches, fewer splits)
};
size_t exact_workspace_bytes(int n, int batch, int kmax);
int launch_exact(const ExactArgs &args, cudaStream_t s);

// synth.cu
int launch_synth_attrs(double *dst, long long rows, int a, long long row_base, double lo, double hi,
                       unsigned long long seed, unsigned long long stream_id, cudaStream_t s);
int launch_synth_labels(int *dst, long long rows, long long row_base, int num_labels, unsigned long long seed,
                        cudaStream_t s);
int launch_synth_k(int *dst, long long rows, long long row_base, int min_k, int max_k,
                   unsigned long long seed, cudaStream_t s);

}  // namespace k
}  // namespace dmlp
