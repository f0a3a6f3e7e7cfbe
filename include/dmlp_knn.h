/* C ABI of libdmlp_knn.so: the sm_100a kernels of the k-NN engine.
 *
 * Every function takes device pointers and a CUDA stream (as void*; NULL is the
 * default stream), launches asynchronously on that stream and returns 0, or a
 * negative value after recording a message for dmlp_last_error().  Nothing here
 * allocates device memory: callers own all buffers, including scratch.
 *
 * Pipeline (one shard of datapoints x one block of queries):
 *   stats -> pack (queries, data) -> dist_topk -> rerank [-> exact for flagged]
 *   -> finalize (merges the lists of all data shards, votes, checksums)
 */
#ifndef DMLP_KNN_C_H_
#define DMLP_KNN_C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char *dmlp_last_error(void);

/* Device properties the planners need: SM count and compute capability. */
int dmlp_device_info(int *sm_count, int *cc_major, int *cc_minor);

/* ---- K7 pack_norms (reference: engine.cpp:79-96, 154-184 repack) -------------
 * A "stats" block is 32 bytes of device memory: {double max(-v), double max(v),
 * float max row norm, float max row error norm}.  The first two fields merge
 * across ranks with one MAX all-reduce over two doubles. */
#define DMLP_STATS_BYTES 32
int dmlp_stats_reset(void *stats, void *stream);
int dmlp_stats_update(const double *src, int64_t count, void *stats, void *stream);
/* src [rows][a] fp64 -> dst [rows][a_pad] fp16 (centred, scaled by a power of two
 * derived from `range`), norm2[rows] = ||x^||^2, errn[rows] >= ||x' - x^||.
 * a_pad must be a multiple of 64.  Row maxima accumulate into `maxima` (may be
 * NULL). */
int dmlp_pack(const double *src, int64_t rows, int a, int a_pad, const void *range, void *dst_fp16,
              float *norm2, float *errn, void *maxima, void *stream);

/* Install a value range computed elsewhere (e.g. by the loader on the host) instead
 * of running dmlp_stats_update over the data.  The row maxima are left untouched. */
int dmlp_stats_set(void *stats, double lo, double hi, void *stream);

/* "decimal6" wire format for the host->device copy (reference: the text grammar's
 * "%.6f" values, generate_input.py:15,20): a value that equals i / 1e6 for a 32-bit
 * integer i travels as i.  encode checks the round trip for every element (sets
 * *bad to 1 otherwise) and is meant for load time; expand reproduces the doubles
 * bit-for-bit on the device. */
int dmlp_encode_decimal6(const double *src, int *dst, int64_t count, int *bad, void *stream);
int dmlp_expand_decimal6(const int *src, double *dst, int64_t count, void *stream);

/* Bit-packed decimal6: when every |i| < 2^(bits-1), a row of `a` integers travels as a little-endian
 * stream of `bits`-bit numbers i + 2^(bits-1), padded to `words` 32-bit words per row (words * 32 >=
 * a * bits).  Values in [-100, 100] need 28 bits: 12.5 % fewer PCIe bytes than int32.  unpack
 * writes either the integers (dst_i32) or the expanded doubles (dst_f64); exactly one is non-NULL. */
int dmlp_pack_bits(const int *src, int64_t rows, int a, int bits, int words, uint32_t *dst, void *stream);
int dmlp_unpack_bits(const uint32_t *src, int64_t rows, int a, int bits, int words, int *dst_i32, double *dst_f64,
                     void *stream);

/* Which k the filter + certified re-rank path takes (everything else is the direct fp64 path,
 * dmlp_exact*): 2 = yes, also for launches that gather operands from peers inside the kernel
 * (kmax <= 96); 1 = yes for launches on resident operands (kmax <= 1024 and about n >= 32 * kmax --
 * one 256-row tile per 8 of k, whole groups of 32 tiles beyond k = 256, where the filter runs as
 * several launches over row ranges of the data, see dmlp_topk_plan); 0 = no. */
int dmlp_fused_supported(int kmax, int64_t n);

/* ---- K1 + K3 dist_topk (reference: engine.cpp:235-256) ------------------------
 * plan_out: {m_blocks, n_tiles, group_m, chunks, k_blocks, grid, kp, cluster}.  `chunks`
 * is the number of candidate lists per query: the candidate buffer holds q * chunks * kp
 * 64-bit keys.  kmax = the largest k any query asks for (0 = unknown): every list keeps at
 * most 32 candidates per query, so larger k is given more lists.  See dmlp_fused_supported
 * for the k the fused path takes.
 * dmlp_topk_plan_ex adds {halves, parts, column chunks per launch, wide, lists per part,
 * re-alignment period, dynamic scheduling (-1 = where there is no re-alignment), 0}: lists =
 * column chunks x halves (two sets of epilogue warps on short rows) x parts (launches over row
 * ranges of the data, k > 256). */
int dmlp_topk_plan(int q, int n, int a_pad, int num_sms, int kmax, int plan_out[8]);
int dmlp_topk_plan_ex(int q, int n, int a_pad, int num_sms, int kmax, int plan_out[16]);
int dmlp_dist_topk(const void *q_fp16, const void *x_fp16, const float *x_norm2, int q, int n, int a_pad,
                   int num_sms, int kmax, uint64_t *cand, void *stream);

/* The same kernel on a slice of the data shard: rows [row_base, row_base + n_part) whose
 * packed rows / norms start at x_fp16_part / x_norm2_part.  The launch writes lists
 * [chunk_base, chunk_base + chunks(q, n_part)) of a candidate buffer that has cand_chunks
 * lists per query, and emits shard-wide row ids.  Lets a caller start filtering while the
 * rest of the shard is still arriving from the host (see DistributedKNN.run_from_host). */
int dmlp_dist_topk_part(const void *q_fp16, const void *x_fp16_part, const float *x_norm2_part, int q, int n_part,
                        int a_pad, int num_sms, int kmax, uint64_t *cand, int cand_chunks, int chunk_base,
                        int row_base, void *stream);

/* ---- K2 exact_rerank (reference: engine.cpp:12-18 semantics) ------------------
 * Writes, per query, kmax slots of (dist, id, label): the exact local top-k in
 * (distance asc, id desc) order, padded with (1e18, -1, -1).  flags[q] != 0 marks
 * a query the filter could not certify; *num_flagged counts them.  Scratch (the
 * survivor pair list) comes from the caller: dmlp_rerank_workspace_bytes(q, kmax) bytes,
 * 256-byte aligned. */
size_t dmlp_rerank_workspace_bytes(int q, int kmax);
int dmlp_rerank(const uint64_t *cand, int chunks, int kp, const double *q_attrs, const double *x_attrs,
                const int *x_labels, const int *k, const float *q_norm2, const float *q_errn,
                const void *range, const void *x_maxima, int q, int n, int a, int a_pad, int id_base, int kmax,
                double *out_dist, int *out_id, int *out_label, int *flags, int *num_flagged, void *workspace,
                size_t workspace_bytes, void *stream);

/* Diagnostics: subsequent dmlp_rerank* launches also record, per query, how many
 * survivors they recomputed in fp64 (device int[q]); NULL turns it off. */
void dmlp_debug_rerank_w(int *w_out);
/* Host replay of the filter kernel's work enumeration for the plan dmlp_topk_plan() returns: 8 ints
 * per item and CTA {cluster, cta rank, row block, chunk, first tile, end tile, wave group, cluster
 * rows in the group}.  Needs no GPU.  Returns the number of items (only max_items are stored). */
int dmlp_debug_topk_items(int q, int n, int a_pad, int num_sms, int kmax, int *out, int max_items);
/* 1 if the plan for this shape uses the wide (256 x 512 per CTA pair) tile. */
int dmlp_debug_topk_wide(int q, int n, int a_pad, int num_sms, int kmax);
/* Diagnostics: with DMLP_TOPK_PROF=1 in the environment the CTA-pair filter kernel records, per CTA,
 * {mma total, mma operand wait, mma accumulator wait, producer total, producer stage wait, producer
 * re-alignment wait, epilogue accumulator wait, epilogue total, epilogue norm staging, accumulator
 * loads, scoring and appends, list compaction, list flush, 3 unused} in SM cycles; this copies the
 * counters of the last launch to host memory (16 values per CTA) and returns the number of CTAs. */
int dmlp_debug_topk_prof(unsigned long long *host_out, int max_ctas);
/* Watchdog of the filter kernel's re-alignment points (CTA pairs that sweep the same data
 * tiles wait for each other so the tiles are shared through L2): a wait that gives up after
 * 2e9 cycles is counted on the device; this reads the count (0 in a healthy run; results are
 * correct either way).  Every dmlp_dist_topk* launch takes its own slice of the per-device
 * counter array, so launches may overlap on different streams. */
int dmlp_topk_sync_timeouts(unsigned *out);

/* ---- exact fp64 brute force (reference: engine.cpp:233-257; bounded per-query lists
 * like the staff baselines' heaps) ---------------------------------
 * Processes `count` queries: rows query_rows[0..count) or, with NULL, rows
 * 0..count, for any k <= n.  Streaming: distances are selected on the fly into
 * per-query lists in the workspace (O(count * kmax) memory, no distance matrix, no
 * library sort).  Scratch comes from the caller, 256-byte aligned:
 * dmlp_exact_workspace_bytes(n, count, kmax) is the recommended size (at most half
 * a gigabyte unless a single query needs more); the launcher works in batches when
 * given less, down to what one query needs. */
size_t dmlp_exact_workspace_bytes(int n, int count, int kmax);
int dmlp_exact(const double *q_attrs, const double *x_attrs, const int *x_labels, const int *k,
               const int *query_rows, int count, int n, int a, int id_base, int kmax, double *out_dist,
               int *out_id, int *out_label, void *workspace, size_t workspace_bytes, void *stream);

/* ---- K5 owner merge + K6 vote/checksum (reference: engine.cpp:289-347) --------
 * dist/id/label: `sources` arrays of [q][kmax], source_stride elements apart, each list
 * sorted by (distance asc, id desc) and padded with (1e18, -1, -1).  Entries are ranked
 * against the other sources' lists by binary search, so any sources * kmax is merged on
 * the device; only sources > 1 with kmax > 2048 needs scratch
 * (dmlp_finalize_workspace_bytes, else 0 / NULL). */
size_t dmlp_finalize_workspace_bytes(int sources, int q, int kmax);
int dmlp_finalize(const double *dist, const int *id, const int *label, int64_t source_stride, int sources,
                  int q, int kmax, const int *k, int *out_label, uint64_t *out_checksum, int *out_id,
                  double *out_dist, void *workspace, size_t workspace_bytes, void *stream);

/* ---- K8 synthetic_fill (reference: generate_input.py:13-21) -------------------
 * stream_id 0 = data attributes, 2 = query attributes. */
int dmlp_synth_attrs(double *dst, int64_t rows, int a, int64_t row_base, double lo, double hi, uint64_t seed,
                     uint64_t stream_id, void *stream);
int dmlp_synth_labels(int *dst, int64_t rows, int64_t row_base, int num_labels, uint64_t seed, void *stream);
int dmlp_synth_k(int *dst, int64_t rows, int64_t row_base, int min_k, int max_k, uint64_t seed, void *stream);

/* ---- whole local pipeline on caller-provided scratch --------------------------
 * stats -> pack -> dist_topk -> rerank for one data shard and one query block.
 * `range` may be NULL (the range is then computed from these two arrays) or a
 * stats block already merged across ranks.  Scratch layout is private; size it
 * with dmlp_knn_local_workspace_bytes. */
size_t dmlp_knn_local_workspace_bytes(int q, int n, int a, int num_sms, int kmax);
int dmlp_knn_local(const double *q_attrs, const int *k, int q, const double *x_attrs, const int *x_labels,
                   int n, int a, int id_base, int kmax, const void *range, void *workspace,
                   size_t workspace_bytes, double *out_dist, int *out_id, int *out_label, int *flags,
                   int *num_flagged, void *stream);

/* ---- symmetric peer memory + device-side signalling (replaces the reference's MPI
 * layer, engine.cpp:33-35, 73-131, 147-209, 282-284) ---------------------------
 * Unlike the kernel entry points above, alloc/open/close/free are set-up calls
 * that do allocate and map memory; call them once, outside the hot path.
 * A flag array is `world` 32-bit words inside a symmetric buffer: word s of rank
 * d's array is written only by rank s. */
int dmlp_symm_alloc(size_t bytes, void **ptr, unsigned char handle_out[64]);   /* cudaMalloc + zero + IPC export */
int dmlp_symm_open(const unsigned char handle[64], void **ptr);                /* map a peer's buffer */
int dmlp_symm_close(void *ptr);
int dmlp_symm_free(void *ptr);
/* peer_flag_ptrs: device array of `world` pointers, entry d = rank d's flag array
 * as mapped here.  Stores `value` into word `my_rank` of every rank's array after
 * a system-scope fence (so all earlier work on this stream is visible first). */
int dmlp_symm_signal(void *peer_flag_ptrs, int world, int my_rank, uint32_t value, void *stream);
/* Spins (on the device, in stream order) until every word of the local flag array
 * has reached `value`; gives up after timeout_seconds and writes 1 + the index of
 * the late rank to *status (device int, may be NULL). */
int dmlp_symm_wait(const void *local_flags, int world, uint32_t value, double timeout_seconds, int *status,
                   void *stream);

/* All-gather of per-rank parts of one region of the symmetric buffers (reference:
 * the row MPI_Bcast that replicates a data shard, engine.cpp:129-131): rank p's
 * part lives at part_off[p] (length part_len[p]) in every rank's buffer; after
 * each rank has signalled word `rank` of the flag array at off_flags with `epoch`,
 * this kernel pulls every other rank's part over NVLink into the local buffer. */
typedef struct dmlp_peer_gather_args {
  int world, rank;
  uint32_t epoch;
  void *peer_base[8];
  uint64_t off_flags;
  uint64_t part_off[8], part_len[8];
  int *status;
  double timeout_seconds;
} dmlp_peer_gather_args;
int dmlp_peer_gather(const dmlp_peer_gather_args *g, void *stream);

/* ---- fused cross-GPU steps ------------------------------------------------------
 * K4, all-gather -> GEMM (reference: the row/column MPI_Bcast of an operand shard,
 * engine.cpp:129-131, 207-209, followed by the distance loop): every rank packs
 * the `slice_rows` queries it owns into its symmetric buffer (at off_qh, rows
 * [rank * slice_rows, ...)), signals pack_flags, and calls dmlp_dist_topk_gather.
 * A communication warp in every GEMM CTA pulls the other ranks' packed rows
 * through the peer mappings just ahead of the tiles that use them, then pulls the
 * fp64 rows (off_q64), norms (off_qn, off_qe) and k (off_k) for the re-rank.
 * rb_count: `world * slice_rows / 128` local counters, zeroed once at set-up;
 * epoch: 1, 2, 3, ... per step.  slice_rows must be a multiple of 256. */
typedef struct dmlp_gather_args {
  int world, rank, slice_rows, a;
  uint32_t epoch;
  void *peer_base[8];
  uint64_t off_qh, off_q64, off_qn, off_qe, off_k, off_pack_flags;
  uint32_t *rb_count;
  int *status;            /* device int: set non-zero if a peer never showed up */
  double timeout_seconds;
} dmlp_gather_args;
int dmlp_dist_topk_gather(const void *x_fp16, const float *x_norm2, int n, int a_pad, int num_sms, int kmax,
                          uint64_t *cand, const dmlp_gather_args *g, void *stream);

/* The same fusion for the data operand (reference: the row MPI_Bcast of a data shard,
 * engine.cpp:129-131): the `world` ranks of a grid row each pack one part of the shard --
 * rows [part_begin[p], part_begin[p+1]) -- into their symmetric buffer (packed rows at
 * off_xh, norms at off_xn / off_xe, fp64 rows at off_x64, labels at off_labels), signal
 * pack_flags and call dmlp_dist_topk_rows.  The communication warps of the GEMM pull the
 * other parts tile by tile in sweep order, then the fp64 rows and labels for the re-rank;
 * afterwards the local buffer holds the whole shard.  tile_count: ceil(n / 256) local
 * counters, zeroed once at set-up.  Run dmlp_maxima_from_norms on the gathered norm
 * arrays before the re-rank. */
typedef struct dmlp_data_gather_args {
  int world, rank, a;
  uint32_t epoch;
  void *peer_base[8];
  uint64_t off_xh, off_xn, off_xe, off_x64, off_labels, off_pack_flags;
  int part_begin[9];
  uint32_t *tile_count;
  int *status;
  double timeout_seconds;
  const struct dmlp_signal_args *done_signal;  /* optional: "I have finished reading your parts", stored by
                                                  the kernel's last block (see dmlp_signal_args below) */
  int x_i32;   /* 1: the region at off_x64 holds decimal6 rows (a int32 per row, a even) instead of fp64 rows */
} dmlp_data_gather_args;
int dmlp_dist_topk_rows(const void *q_fp16, int q, int n, int a_pad, int num_sms, int kmax, uint64_t *cand,
                        const dmlp_data_gather_args *g, void *stream);
int dmlp_maxima_from_norms(const float *norm2, const float *errn, int64_t rows, void *maxima, void *stream);

/* K5, GEMM/top-k -> reduce-scatter of k-lists (reference: MPI_Gather of k-lists to
 * row 0 + re-select, engine.cpp:283-308): dmlp_rerank with its outputs redirected
 * into the owner's receive buffers [world][slice_rows][kmax] (off_dist, off_id,
 * off_label inside the owner's symmetric buffer), written through NVLink from the
 * kernel.  Follow with dmlp_symm_signal; the owner runs dmlp_symm_wait and
 * dmlp_finalize over its receive buffers. */
typedef struct dmlp_scatter_args {
  int world, rank, slice_rows;
  void *peer_base[8];
  uint64_t off_dist, off_id, off_label;
} dmlp_scatter_args;
int dmlp_rerank_scatter(const uint64_t *cand, int chunks, int kp, const double *q_attrs, const double *x_attrs,
                        const int *x_labels, const int *k, const float *q_norm2, const float *q_errn,
                        const void *range, const void *x_maxima, int q, int n, int a, int a_pad, int id_base,
                        int kmax, int *flags, int *num_flagged, const dmlp_scatter_args *sc, void *workspace,
                        size_t workspace_bytes, void *stream);

/* ---- struct-based forms of the re-rank and of the repair pass -----------------------
 * dmlp_rerank_ex is dmlp_rerank / dmlp_rerank_scatter with the optional extras:
 *   final_label / final_checksum  (single data shard only, scatter == NULL): the emit kernel
 *       also votes and folds the checksum, so no dmlp_finalize launch is needed (K6 fused
 *       into K2; reference: engine.cpp:320-347, common.cpp:59-68);
 *   flagged_rows / flagged_cap    the indices of the queries the filter could not certify,
 *       in the order they bumped *num_flagged (entries beyond flagged_cap are dropped);
 *   signal                        K5: when the whole launch has finished, store `value` into
 *       this rank's word of every peer's flag array (release at system scope), from inside
 *       the kernel -- no separate signal launch.
 * dmlp_repair_flagged re-runs the first min(*num_flagged, cap) of those rows on the direct
 * fp64 kernel and overwrites their lists (locally or, with scatter, in the owners' receive
 * buffers); with final_label set it also re-finalizes those rows.  The count stays on the
 * device: the pass is enqueued unconditionally and does nothing when no query was flagged,
 * so a step needs no host synchronisation in the middle.  If *num_flagged > cap the caller
 * must handle the rest (check the counter once the step is over).
 * Workspace: dmlp_exact_workspace_bytes(n, cap, kmax). */
/* dmlp_pack with the two ends of a flag protocol folded in: the kernel first waits (wait, may be
 * NULL) until every peer has finished reading the rows it is about to overwrite, and its last
 * block announces the new rows (signal, may be NULL).  rows == 0 still signals. */
struct dmlp_signal_args;
struct dmlp_wait_args;
int dmlp_pack_ex(const double *src, int64_t rows, int a, int a_pad, const void *range, void *dst_fp16, float *norm2,
                 float *errn, void *maxima, const struct dmlp_wait_args *wait, const struct dmlp_signal_args *signal,
                 void *stream);
/* The same with the source rows in decimal6 form: int32 i per value, value = i / 1e6 (the loader's
 * wire format kept resident).  The kernel rebuilds the parser's double exactly (one multiply and a
 * two-FMA correction step, checked against IEEE division for every 32-bit i) and reads half the
 * bytes.  wait / signal may be NULL. */
int dmlp_pack_i32(const int *src, int64_t rows, int a, int a_pad, const void *range, void *dst_fp16, float *norm2,
                  float *errn, void *maxima, const struct dmlp_wait_args *wait, const struct dmlp_signal_args *signal,
                  void *stream);
/* cudaMemsetAsync(ptr, 0, bytes) on the stream (counters, flags) without a library kernel. */
int dmlp_zero(void *ptr, size_t bytes, void *stream);

typedef struct dmlp_signal_args {
  void *peer_flags[8];     /* every rank's flag array as mapped here; word `rank` is written */
  int world, rank;
  uint32_t value;
  uint32_t *done_counter;  /* local device word, zero at set-up; counts finished blocks and resets itself */
  uint32_t *mc_flags;      /* NULL, or the flag array through an NVSwitch multicast mapping that covers every
                              rank's array (cuMulticast*): the signal is then ONE multimem store that the
                              switch replicates, instead of `world` stores; peer_flags is not used */
} dmlp_signal_args;

/* The receiving side of `signal`: the kernel spins (acquire, system scope) until every one of the
 * `world` words of the local flag array has reached `value` before it touches what the peers
 * wrote; after timeout_seconds it writes 600 + the late rank to *status and carries on. */
typedef struct dmlp_wait_args {
  const uint32_t *flags;
  int world;
  uint32_t value;
  double timeout_seconds;
  int *status;
} dmlp_wait_args;

typedef struct dmlp_rerank_args {
  const uint64_t *cand;
  int chunks, kp;
  const double *q_attrs, *x_attrs;
  const int *x_labels, *k;
  const float *q_norm2, *q_errn;
  const void *range, *x_maxima;
  int q, n, a, a_pad, id_base, kmax;
  double *out_dist;
  int *out_id, *out_label;
  int *flags, *num_flagged;
  int *flagged_rows;
  int flagged_cap;
  int *final_label;
  uint64_t *final_checksum;
  const dmlp_scatter_args *scatter;
  const dmlp_signal_args *signal;
  void *workspace;
  size_t workspace_bytes;
  const int *x_attrs_i32;   /* optional: the data rows in decimal6 form (int32 i, value i / 1e6), read
                               instead of x_attrs: half the bytes of the exact pass.  Needs an even a. */
} dmlp_rerank_args;
int dmlp_rerank_ex(const dmlp_rerank_args *r, void *stream);

typedef struct dmlp_repair_args {
  const double *q_attrs, *x_attrs;
  const int *x_labels, *k;
  const int *flagged_rows, *num_flagged;
  int cap;
  int q, n, a, id_base, kmax;
  double *out_dist;
  int *out_id, *out_label;
  int *final_label;
  uint64_t *final_checksum;
  const dmlp_scatter_args *scatter;
  const dmlp_signal_args *signal;
  void *workspace;
  size_t workspace_bytes;
  const int *x_attrs_i32;   /* optional, as in dmlp_rerank_args */
} dmlp_repair_args;
int dmlp_repair_flagged(const dmlp_repair_args *r, void *stream);

/* dmlp_exact with the K5 extras: `scatter` sends the lists into the owners' receive buffers (the
 * query rows are then indices in the padded column space, owner = row / slice_rows), `signal`
 * announces them from the kernel's last block. */
typedef struct dmlp_exact_args {
  const double *q_attrs, *x_attrs;
  const int *x_labels, *k, *query_rows;
  int count, n, a, id_base, kmax;
  double *out_dist;
  int *out_id, *out_label;
  const dmlp_scatter_args *scatter;
  const dmlp_signal_args *signal;
  void *workspace;
  size_t workspace_bytes;
  const int *x_attrs_i32;   /* optional, as in dmlp_rerank_args */
} dmlp_exact_args;
int dmlp_exact_ex(const dmlp_exact_args *g, void *stream);

/* dmlp_finalize with the extras: `wait` (K5 owner side: the lists come from the peers, see
 * dmlp_signal_args; no separate wait launch), and a row subset whose length may live on the
 * device (query_rows / count / count_dev, e.g. the repaired rows of this step). */
typedef struct dmlp_finalize_args {
  const double *dist;
  const int *id, *label;
  int64_t source_stride;
  int sources, q, kmax;
  const int *k;
  int *out_label;
  uint64_t *out_checksum;
  int *out_id;
  double *out_dist;
  const int *query_rows;
  int count;
  const int *count_dev;
  const dmlp_wait_args *wait;
  void *workspace;
  size_t workspace_bytes;
} dmlp_finalize_args;
int dmlp_finalize_ex(const dmlp_finalize_args *f, void *stream);

/* ---- the same pipeline for callers that cannot allocate -------------------------
 * (No counterpart in the reference: it has no GPU code and no scratch buffers.)
 * Same inputs, outputs and results as dmlp_knn_local with range == NULL, but the
 * scratch is a fixed DMLP_STATIC_SCRATCH_BYTES array built into the library, so
 * the call needs no workspace and performs no cudaMalloc/cudaFree or any other
 * allocating call.  It exists for hosts that must not allocate once running:
 * plug-ins inside another program's CUDA context, real-time or certified loops
 * where allocation latency and failure are unacceptable, processes confined to a
 * memory reservation they do not control.
 *  - Limit: problems whose dmlp_knn_local_workspace_bytes() exceeds
 *    DMLP_STATIC_SCRATCH_BYTES are refused (-3, message in dmlp_last_error()).
 *  - Every GPU that loads the library keeps the array reserved until exit.
 *  - The first call in each CUDA context (by default: one per device; a host
 *    application's own context counts separately) looks up that context's address
 *    of the array; make that call at start-up, before any stream capture there.
 *  - An empty data shard (n == 0) is answered with padding lists without touching
 *    the scratch.
 *  - All calls share the one array, which keeps the previous call's contents;
 *    the call re-initialises what it accumulates into.  Calls that overlap on
 *    the same GPU (different streams or threads) race on it: serialise them. */
#define DMLP_STATIC_SCRATCH_BYTES (64u << 20)
int dmlp_knn_local_static(const double *q_attrs, const int *k, int q, const double *x_attrs, const int *x_labels,
                          int n, int a, int id_base, int kmax, double *out_dist, int *out_id, int *out_label,
                          int *flags, int *num_flagged, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DMLP_KNN_C_H_ */
