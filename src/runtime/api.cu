// C ABI over the kernel launchers.  See include/dmlp_knn.h.
#include "dmlp_knn.h"

#include <string>

#include "../kernels/common.cuh"
#include "../kernels/kernels.h"

namespace dmlp {
namespace k {
namespace {
thread_local std::string g_error;
}
void set_error(const std::string &msg) { g_error = msg; }
const char *get_error() { return g_error.c_str(); }
}  // namespace k
}  // namespace dmlp

using namespace dmlp::k;

static_assert(sizeof(PackStats) <= DMLP_STATS_BYTES, "stats block size");

static __global__ void stats_set_kernel(PackStats *st, double neg_lo, double hi) {
  if (threadIdx.x == 0) { st->neg_lo = neg_lo; st->hi = hi; }
}

// An empty shard has no neighbours to offer: every list is padding, no query is flagged.
static __global__ void __launch_bounds__(256) fill_empty_lists_kernel(double *dist, int *id, int *label, int *flags,
                                                               long long slots, int q) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < slots; i += stride) {
    dist[i] = 1e18;
    id[i] = -1;
    label[i] = -1;
    if (i < q) flags[i] = 0;
  }
}
namespace {
inline cudaStream_t st(void *s) { return static_cast<cudaStream_t>(s); }
inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }
inline int pad64(int a) { return (a + 63) / 64 * 64; }

int fill_scatter(const dmlp_scatter_args *sc, ScatterArgs *out) {
  DMLP_REQUIRE(sc->world >= 1 && sc->world <= kMaxPeers && sc->slice_rows > 0, "scatter: world must be in [1, 8]");
  out->world = sc->world; out->rank = sc->rank; out->slice_rows = sc->slice_rows;
  for (int i = 0; i < sc->world; ++i) out->peer_base[i] = static_cast<unsigned char *>(sc->peer_base[i]);
  out->off_dist = sc->off_dist; out->off_id = sc->off_id; out->off_label = sc->off_label;
  return 0;
}

int fill_signal(const dmlp_signal_args *sg, SignalArgs *out) {
  if (sg == nullptr || sg->world <= 0) return 0;
  DMLP_REQUIRE(sg->world <= kMaxPeers && sg->done_counter != nullptr, "signal: world must be in [1, 8] and done_counter set");
  out->world = sg->world; out->rank = sg->rank; out->value = sg->value; out->done_counter = sg->done_counter;
  out->mc_flags = sg->mc_flags;
  for (int i = 0; i < sg->world; ++i) out->peer_flags[i] = static_cast<unsigned *>(sg->peer_flags[i]);
  return 0;
}

int fill_wait(const dmlp_wait_args *w, WaitArgs *out) {
  if (w == nullptr || w->world <= 0) return 0;
  DMLP_REQUIRE(w->world <= kMaxPeers && w->flags != nullptr, "wait: world must be in [1, 8] and flags set");
  out->flags = w->flags; out->world = w->world; out->value = w->value; out->status = w->status;
  out->timeout_cycles = static_cast<long long>((w->timeout_seconds > 0 ? w->timeout_seconds : 20.0) * 1.5e9);
  return 0;
}

int current_sm_count(int *out) {
  int dev = 0;
  DMLP_CUDA_OK(cudaGetDevice(&dev));
  DMLP_CUDA_OK(cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev));
  return 0;
}
}  // namespace

extern "C" {

const char *dmlp_last_error(void) { return get_error(); }

int dmlp_device_info(int *sm_count, int *cc_major, int *cc_minor) {
  int dev = 0;
  DMLP_CUDA_OK(cudaGetDevice(&dev));
  DMLP_CUDA_OK(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
  DMLP_CUDA_OK(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  DMLP_CUDA_OK(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
  return 0;
}

int dmlp_stats_reset(void *stats, void *stream) {
  return launch_stats_reset(static_cast<PackStats *>(stats), st(stream));
}

int dmlp_stats_update(const double *src, int64_t count, void *stats, void *stream) {
  return launch_stats(src, count, static_cast<PackStats *>(stats), st(stream));
}

int dmlp_pack(const double *src, int64_t rows, int a, int a_pad, const void *range, void *dst_fp16, float *norm2,
              float *errn, void *maxima, void *stream) {
  DMLP_REQUIRE(a_pad % 64 == 0, "pack: a_pad must be a multiple of 64");
  return launch_pack(src, rows, a, a_pad, static_cast<const PackStats *>(range), static_cast<__half *>(dst_fp16),
                     norm2, errn, static_cast<PackStats *>(maxima), st(stream));
}

int dmlp_pack_ex(const double *src, int64_t rows, int a, int a_pad, const void *range, void *dst_fp16, float *norm2,
                 float *errn, void *maxima, const dmlp_wait_args *wait, const dmlp_signal_args *signal, void *stream) {
  DMLP_REQUIRE(a_pad % 64 == 0, "pack: a_pad must be a multiple of 64");
  WaitArgs wa;
  SignalArgs sg;
  if (int rc = fill_wait(wait, &wa)) return rc;
  if (int rc = fill_signal(signal, &sg)) return rc;
  return launch_pack(src, rows, a, a_pad, static_cast<const PackStats *>(range), static_cast<__half *>(dst_fp16),
                     norm2, errn, static_cast<PackStats *>(maxima), st(stream), &wa, &sg);
}

int dmlp_pack_i32(const int *src, int64_t rows, int a, int a_pad, const void *range, void *dst_fp16, float *norm2,
                  float *errn, void *maxima, const dmlp_wait_args *wait, const dmlp_signal_args *signal, void *stream) {
  DMLP_REQUIRE(a_pad % 64 == 0, "pack: a_pad must be a multiple of 64");
  WaitArgs wa;
  SignalArgs sg;
  if (int rc = fill_wait(wait, &wa)) return rc;
  if (int rc = fill_signal(signal, &sg)) return rc;
  return launch_pack(src, rows, a, a_pad, static_cast<const PackStats *>(range), static_cast<__half *>(dst_fp16),
                     norm2, errn, static_cast<PackStats *>(maxima), st(stream), &wa, &sg, /*src_i32=*/true);
}

int dmlp_zero(void *ptr, size_t bytes, void *stream) {
  DMLP_CUDA_OK(cudaMemsetAsync(ptr, 0, bytes, st(stream)));
  return 0;
}

int dmlp_expand_decimal6(const int *src, double *dst, int64_t count, void *stream) {
  return launch_expand_decimal6(src, dst, count, st(stream));
}

int dmlp_pack_bits(const int *src, int64_t rows, int a, int bits, int words, uint32_t *dst, void *stream) {
  return launch_pack_bits(src, rows, a, bits, words, dst, st(stream));
}

int dmlp_unpack_bits(const uint32_t *src, int64_t rows, int a, int bits, int words, int *dst_i32, double *dst_f64,
                     void *stream) {
  return launch_unpack_bits(src, rows, a, bits, words, dst_i32, dst_f64, st(stream));
}

int dmlp_encode_decimal6(const double *src, int *dst, int64_t count, int *bad, void *stream) {
  return launch_encode_decimal6(src, dst, count, bad, st(stream));
}

int dmlp_stats_set(void *stats, double lo, double hi, void *stream) {
  // a kernel, not a copy: a host-to-device copy would queue behind whatever large transfers the
  // copy engine is busy with (the end-to-end loop keeps it busy all the time)
  stats_set_kernel<<<1, 32, 0, st(stream)>>>(static_cast<PackStats *>(stats), -lo, hi);
  DMLP_CUDA_OK(cudaGetLastError());
  return 0;
}

int dmlp_topk_plan(int q, int n, int a_pad, int num_sms, int kmax, int plan_out[8]) {
  TopkPlan p = make_topk_plan(q, n, a_pad, 32, num_sms, kmax);
  plan_out[0] = p.m_blocks; plan_out[1] = p.n_tiles; plan_out[2] = p.group_m; plan_out[3] = p.lists;
  plan_out[4] = p.k_blocks; plan_out[5] = p.grid; plan_out[6] = p.kp; plan_out[7] = p.cluster;
  return 0;
}

int dmlp_fused_supported(int kmax, int64_t n) {
  if (kmax < 1 || n < 1) return 0;
  if (kmax <= 96) return 2;
  if (kmax > 1024) return 0;
  // beyond 96: one 32-key list per 8 of k and a list needs a data tile (256 rows) of its own; a
  // launch has at most 32 lists, so beyond k = 256 the filter runs as several launches over row
  // ranges of at least 32 tiles each (make_topk_plan) -- all in all n >= about 32 k
  const int64_t n_tiles = (n + 255) / 256;
  const int c_need = (kmax + 7) / 8, parts = (c_need + 31) / 32;
  return (parts > 1 ? n_tiles >= 32LL * parts : n_tiles >= c_need) ? 1 : 0;
}

int dmlp_topk_plan_ex(int q, int n, int a_pad, int num_sms, int kmax, int plan_out[16]) {
  TopkPlan p = make_topk_plan(q, n, a_pad, 32, num_sms, kmax);
  for (int i = 0; i < 16; ++i) plan_out[i] = 0;
  if (int rc = dmlp_topk_plan(q, n, a_pad, num_sms, kmax, plan_out)) return rc;
  plan_out[8] = p.halves; plan_out[9] = p.parts; plan_out[10] = p.chunks; plan_out[11] = p.wide;
  plan_out[12] = p.part_chunks; plan_out[13] = p.sync; plan_out[14] = p.dyn;
  return 0;
}

int dmlp_debug_topk_wide(int q, int n, int a_pad, int num_sms, int kmax) {
  return make_topk_plan(q, n, a_pad, 32, num_sms, kmax).wide;
}

int dmlp_debug_topk_items(int q, int n, int a_pad, int num_sms, int kmax, int *out, int max_items) {
  TopkPlan p = make_topk_plan(q, n, a_pad, 32, num_sms, kmax);
  return enumerate_topk_items(p, out, max_items);
}

int dmlp_dist_topk(const void *q_fp16, const void *x_fp16, const float *x_norm2, int q, int n, int a_pad,
                   int num_sms, int kmax, uint64_t *cand, void *stream) {
  if (num_sms <= 0)
    if (int rc = current_sm_count(&num_sms)) return rc;
  TopkPlan p = make_topk_plan(q, n, a_pad, 32, num_sms, kmax);
  return launch_dist_topk(static_cast<const __half *>(q_fp16), static_cast<const __half *>(x_fp16), x_norm2, p,
                          reinterpret_cast<unsigned long long *>(cand), nullptr, st(stream));
}

int dmlp_rerank_ex(const dmlp_rerank_args *g, void *stream) {
  DMLP_REQUIRE(g != nullptr, "rerank: null arguments");
  RerankArgs r;
  r.workspace = g->workspace; r.workspace_bytes = g->workspace_bytes;
  r.cand = reinterpret_cast<const unsigned long long *>(g->cand);
  r.chunks = g->chunks; r.kp = g->kp;
  r.q_attrs = g->q_attrs; r.x_attrs = g->x_attrs; r.x_labels = g->x_labels; r.k = g->k;
  r.q_norm2 = g->q_norm2; r.q_errn = g->q_errn;
  r.range = static_cast<const PackStats *>(g->range);
  r.x_maxima = static_cast<const PackStats *>(g->x_maxima);
  r.q = g->q; r.n = g->n; r.a = g->a; r.a_pad = g->a_pad; r.id_base = g->id_base; r.kmax = g->kmax;
  r.out_dist = g->out_dist; r.out_id = g->out_id; r.out_label = g->out_label;
  r.flags = g->flags; r.num_flagged = g->num_flagged;
  r.flagged_rows = g->flagged_rows; r.flagged_cap = g->flagged_rows ? g->flagged_cap : 0;
  r.final_label = g->final_label;
  r.final_checksum = reinterpret_cast<unsigned long long *>(g->final_checksum);
  r.x_attrs_i32 = g->x_attrs_i32;
  if (g->scatter != nullptr) {
    const dmlp_scatter_args *sc = g->scatter;
    DMLP_REQUIRE(sc->world >= 2 && sc->world <= kMaxPeers, "rerank: scatter world must be in [2, 8]");
    DMLP_REQUIRE(g->q == sc->world * sc->slice_rows, "rerank: with scatter, q must equal world * slice_rows");
    DMLP_REQUIRE(g->final_label == nullptr, "rerank: the folded vote needs a single data shard (no scatter)");
    if (int rc = fill_scatter(sc, &r.scatter)) return rc;
  } else {
    DMLP_REQUIRE(g->out_dist && g->out_id && g->out_label, "rerank: output lists missing");
  }
  DMLP_REQUIRE((g->final_label == nullptr) == (g->final_checksum == nullptr),
               "rerank: final_label and final_checksum come together");
  if (int rc = fill_signal(g->signal, &r.signal)) return rc;
  return launch_rerank(r, st(stream));
}

int dmlp_rerank(const uint64_t *cand, int chunks, int kp, const double *q_attrs, const double *x_attrs,
                const int *x_labels, const int *k, const float *q_norm2, const float *q_errn, const void *range,
                const void *x_maxima, int q, int n, int a, int a_pad, int id_base, int kmax, double *out_dist,
                int *out_id, int *out_label, int *flags, int *num_flagged, void *workspace, size_t workspace_bytes,
                void *stream) {
  dmlp_rerank_args g = {};
  g.cand = cand; g.chunks = chunks; g.kp = kp; g.q_attrs = q_attrs; g.x_attrs = x_attrs; g.x_labels = x_labels;
  g.k = k; g.q_norm2 = q_norm2; g.q_errn = q_errn; g.range = range; g.x_maxima = x_maxima;
  g.q = q; g.n = n; g.a = a; g.a_pad = a_pad; g.id_base = id_base; g.kmax = kmax;
  g.out_dist = out_dist; g.out_id = out_id; g.out_label = out_label; g.flags = flags; g.num_flagged = num_flagged;
  g.workspace = workspace; g.workspace_bytes = workspace_bytes;
  return dmlp_rerank_ex(&g, stream);
}

int dmlp_dist_topk_gather(const void *x_fp16, const float *x_norm2, int n, int a_pad, int num_sms, int kmax,
                          uint64_t *cand, const dmlp_gather_args *g, void *stream) {
  DMLP_REQUIRE(g != nullptr && g->world >= 1 && g->world <= kMaxPeers, "dist_topk_gather: world must be in [1, 8]");
  DMLP_REQUIRE(g->slice_rows > 0 && g->slice_rows % 256 == 0, "dist_topk_gather: slice_rows must be a multiple of 256");
  if (num_sms <= 0)
    if (int rc = current_sm_count(&num_sms)) return rc;
  const int q = g->world * g->slice_rows;
  TopkPlan p = make_topk_plan(q, n, a_pad, 32, num_sms, kmax);
  DMLP_REQUIRE(p.pair == 1, "dist_topk_gather: needs the CTA-pair kernel");
  // start the sweep at this rank's own slice: its rows are ready without any transfer,
  // and every GPU pulls from a different peer first
  p.rot = static_cast<int>((static_cast<long long>(g->rank) * g->slice_rows / 128 / p.cluster) % p.mc_blocks);
  GatherArgs ga;
  ga.world = g->world; ga.rank = g->rank; ga.slice_rows = g->slice_rows; ga.a = g->a; ga.epoch = g->epoch;
  for (int i = 0; i < g->world; ++i) ga.peer_base[i] = static_cast<unsigned char *>(g->peer_base[i]);
  ga.off_qh = g->off_qh; ga.off_q64 = g->off_q64; ga.off_qn = g->off_qn; ga.off_qe = g->off_qe; ga.off_k = g->off_k;
  ga.off_pack_flags = g->off_pack_flags;
  ga.rb_count = g->rb_count; ga.status = g->status;
  ga.timeout_cycles = static_cast<long long>((g->timeout_seconds > 0 ? g->timeout_seconds : 20.0) * 1.5e9);
  const __half *qh = reinterpret_cast<const __half *>(ga.peer_base[ga.rank] + ga.off_qh);
  return launch_dist_topk(qh, static_cast<const __half *>(x_fp16), x_norm2, p,
                          reinterpret_cast<unsigned long long *>(cand), &ga, st(stream));
}

int dmlp_dist_topk_part(const void *q_fp16, const void *x_fp16_part, const float *x_norm2_part, int q, int n_part,
                        int a_pad, int num_sms, int kmax, uint64_t *cand, int cand_chunks, int chunk_base,
                        int row_base, void *stream) {
  if (num_sms <= 0)
    if (int rc = current_sm_count(&num_sms)) return rc;
  TopkPlan p = make_topk_plan(q, n_part, a_pad, 32, num_sms, kmax);
  DMLP_REQUIRE(chunk_base >= 0 && chunk_base + p.lists <= cand_chunks,
               "dist_topk_part: this launch's lists do not fit the candidate buffer");
  p.cand_chunks = cand_chunks;
  p.chunk_base = chunk_base;
  p.row_base = row_base;
  return launch_dist_topk(static_cast<const __half *>(q_fp16), static_cast<const __half *>(x_fp16_part), x_norm2_part,
                          p, reinterpret_cast<unsigned long long *>(cand), nullptr, st(stream));
}

int dmlp_dist_topk_rows(const void *q_fp16, int q, int n, int a_pad, int num_sms, int kmax, uint64_t *cand,
                        const dmlp_data_gather_args *g, void *stream) {
  DMLP_REQUIRE(g != nullptr && g->world >= 1 && g->world <= kMaxPeers, "dist_topk_rows: world must be in [1, 8]");
  if (num_sms <= 0)
    if (int rc = current_sm_count(&num_sms)) return rc;
  TopkPlan p = make_topk_plan(q, n, a_pad, 32, num_sms, kmax);
  DMLP_REQUIRE(p.pair == 1 || g->world == 1, "dist_topk_rows: needs the CTA-pair kernel (at least 256 queries)");
  DataGatherArgs dg;
  dg.world = g->world; dg.rank = g->rank; dg.a = g->a; dg.epoch = g->epoch;
  for (int i = 0; i < g->world; ++i) dg.peer_base[i] = static_cast<unsigned char *>(g->peer_base[i]);
  for (int i = 0; i <= g->world; ++i) dg.part_begin[i] = g->part_begin[i];
  DMLP_REQUIRE(dg.part_begin[0] == 0 && dg.part_begin[g->world] == n, "dist_topk_rows: parts must cover [0, n)");
  dg.off_xh = g->off_xh; dg.off_xn = g->off_xn; dg.off_xe = g->off_xe; dg.off_x64 = g->off_x64;
  dg.off_labels = g->off_labels; dg.off_pack_flags = g->off_pack_flags;
  dg.tile_count = g->tile_count; dg.status = g->status;
  dg.timeout_cycles = static_cast<long long>((g->timeout_seconds > 0 ? g->timeout_seconds : 20.0) * 1.5e9);
  if (int rc = fill_signal(g->done_signal, &dg.done)) return rc;
  dg.x_units = g->x_i32 ? g->a / 2 : g->a;
  DMLP_REQUIRE(!g->x_i32 || g->a % 2 == 0, "dist_topk_rows: decimal6 rows need an even row length");
  const __half *xh = reinterpret_cast<const __half *>(dg.peer_base[dg.rank] + dg.off_xh);
  const float *xn = reinterpret_cast<const float *>(dg.peer_base[dg.rank] + dg.off_xn);
  return launch_dist_topk(static_cast<const __half *>(q_fp16), xh, xn, p,
                          reinterpret_cast<unsigned long long *>(cand), nullptr, st(stream), &dg);
}

int dmlp_maxima_from_norms(const float *norm2, const float *errn, int64_t rows, void *maxima, void *stream) {
  return launch_maxima_from_norms(norm2, errn, rows, static_cast<PackStats *>(maxima), st(stream));
}

int dmlp_rerank_scatter(const uint64_t *cand, int chunks, int kp, const double *q_attrs, const double *x_attrs,
                        const int *x_labels, const int *k, const float *q_norm2, const float *q_errn,
                        const void *range, const void *x_maxima, int q, int n, int a, int a_pad, int id_base,
                        int kmax, int *flags, int *num_flagged, const dmlp_scatter_args *sc, void *workspace,
                        size_t workspace_bytes, void *stream) {
  DMLP_REQUIRE(sc != nullptr, "rerank_scatter: scatter arguments missing");
  dmlp_rerank_args g = {};
  g.cand = cand; g.chunks = chunks; g.kp = kp; g.q_attrs = q_attrs; g.x_attrs = x_attrs; g.x_labels = x_labels;
  g.k = k; g.q_norm2 = q_norm2; g.q_errn = q_errn; g.range = range; g.x_maxima = x_maxima;
  g.q = q; g.n = n; g.a = a; g.a_pad = a_pad; g.id_base = id_base; g.kmax = kmax;
  g.flags = flags; g.num_flagged = num_flagged; g.scatter = sc;
  g.workspace = workspace; g.workspace_bytes = workspace_bytes;
  return dmlp_rerank_ex(&g, stream);
}

int dmlp_repair_flagged(const dmlp_repair_args *g, void *stream) {
  DMLP_REQUIRE(g != nullptr && g->cap > 0 && g->flagged_rows != nullptr && g->num_flagged != nullptr,
               "repair_flagged: needs the flagged-row list, its device counter and cap > 0");
  ExactArgs e;
  e.q_attrs = g->q_attrs; e.x_attrs = g->x_attrs; e.x_labels = g->x_labels; e.k = g->k;
  e.query_rows = g->flagged_rows; e.count = g->cap; e.count_dev = g->num_flagged;
  e.x_attrs_i32 = g->x_attrs_i32;
  e.n = g->n; e.a = g->a; e.id_base = g->id_base; e.kmax = g->kmax;
  e.out_dist = g->out_dist; e.out_id = g->out_id; e.out_label = g->out_label;
  e.workspace = g->workspace; e.workspace_bytes = g->workspace_bytes;
  if (g->scatter != nullptr) {
    DMLP_REQUIRE(g->final_label == nullptr, "repair_flagged: re-finalizing needs a single data shard (no scatter)");
    if (int rc = fill_scatter(g->scatter, &e.scatter)) return rc;
  } else {
    DMLP_REQUIRE(g->out_dist && g->out_id && g->out_label, "repair_flagged: output lists missing");
  }
  if (int rc = fill_signal(g->signal, &e.signal)) return rc;
  if (int rc = launch_exact(e, st(stream))) return rc;
  if (g->final_label != nullptr) {
    FinalizeArgs f;
    f.dist = g->out_dist; f.id = g->out_id; f.label = g->out_label;
    f.source_stride = static_cast<long long>(g->q) * g->kmax;
    f.sources = 1; f.q = g->q; f.kmax = g->kmax; f.k = g->k;
    f.out_label = g->final_label; f.out_checksum = reinterpret_cast<unsigned long long *>(g->final_checksum);
    f.out_id = nullptr; f.out_dist = nullptr;
    f.query_rows = g->flagged_rows; f.count = g->cap; f.count_dev = g->num_flagged;
    return launch_finalize(f, st(stream));
  }
  return 0;
}

void dmlp_debug_rerank_w(int *w_out) { set_rerank_debug_w(w_out); }
int dmlp_debug_topk_prof(unsigned long long *host_out, int max_ctas) { return dist_topk_prof_read(host_out, max_ctas); }
int dmlp_topk_sync_timeouts(unsigned *out) { return dist_topk_sync_timeouts(out); }

size_t dmlp_rerank_workspace_bytes(int q, int kmax) { return rerank_workspace_bytes(q, kmax); }

size_t dmlp_exact_workspace_bytes(int n, int count, int kmax) {
  // everything at once if that stays modest, else half a gigabyte (the launcher then works in
  // batches), and never less than one query needs
  const size_t full = exact_workspace_bytes(n, count, kmax), one = exact_workspace_bytes(n, 1, kmax);
  const size_t cap = size_t(512) << 20;
  return full <= cap ? full : (one > cap ? one : cap);
}

int dmlp_exact(const double *q_attrs, const double *x_attrs, const int *x_labels, const int *k,
               const int *query_rows, int count, int n, int a, int id_base, int kmax, double *out_dist,
               int *out_id, int *out_label, void *workspace, size_t workspace_bytes, void *stream) {
  ExactArgs e;
  e.q_attrs = q_attrs; e.x_attrs = x_attrs; e.x_labels = x_labels; e.k = k; e.query_rows = query_rows;
  e.count = count; e.n = n; e.a = a; e.id_base = id_base; e.kmax = kmax;
  e.out_dist = out_dist; e.out_id = out_id; e.out_label = out_label;
  e.workspace = workspace; e.workspace_bytes = workspace_bytes;
  return launch_exact(e, st(stream));
}

int dmlp_exact_ex(const dmlp_exact_args *g, void *stream) {
  DMLP_REQUIRE(g != nullptr, "exact: null arguments");
  ExactArgs e;
  e.q_attrs = g->q_attrs; e.x_attrs = g->x_attrs; e.x_labels = g->x_labels; e.k = g->k; e.query_rows = g->query_rows;
  e.count = g->count; e.n = g->n; e.a = g->a; e.id_base = g->id_base; e.kmax = g->kmax;
  e.out_dist = g->out_dist; e.out_id = g->out_id; e.out_label = g->out_label;
  e.workspace = g->workspace; e.workspace_bytes = g->workspace_bytes;
  e.x_attrs_i32 = g->x_attrs_i32;
  if (g->scatter != nullptr) {
    if (int rc = fill_scatter(g->scatter, &e.scatter)) return rc;
  } else {
    DMLP_REQUIRE(g->out_dist && g->out_id && g->out_label, "exact: output lists missing");
  }
  if (int rc = fill_signal(g->signal, &e.signal)) return rc;
  return launch_exact(e, st(stream));
}

size_t dmlp_finalize_workspace_bytes(int sources, int q, int kmax) { return finalize_workspace_bytes(sources, q, kmax); }

int dmlp_finalize_ex(const dmlp_finalize_args *g, void *stream) {
  DMLP_REQUIRE(g != nullptr, "finalize: null arguments");
  FinalizeArgs f;
  f.dist = g->dist; f.id = g->id; f.label = g->label; f.source_stride = g->source_stride;
  f.sources = g->sources; f.q = g->q; f.kmax = g->kmax; f.k = g->k;
  f.out_label = g->out_label; f.out_checksum = reinterpret_cast<unsigned long long *>(g->out_checksum);
  f.out_id = g->out_id; f.out_dist = g->out_dist;
  f.workspace = g->workspace; f.workspace_bytes = g->workspace_bytes;
  f.query_rows = g->query_rows; f.count = g->count; f.count_dev = g->count_dev;
  if (int rc = fill_wait(g->wait, &f.wait)) return rc;
  return launch_finalize(f, st(stream));
}

int dmlp_finalize(const double *dist, const int *id, const int *label, int64_t source_stride, int sources, int q,
                  int kmax, const int *k, int *out_label, uint64_t *out_checksum, int *out_id, double *out_dist,
                  void *workspace, size_t workspace_bytes, void *stream) {
  FinalizeArgs f;
  f.workspace = workspace; f.workspace_bytes = workspace_bytes;
  f.dist = dist; f.id = id; f.label = label; f.source_stride = source_stride;
  f.sources = sources; f.q = q; f.kmax = kmax; f.k = k;
  f.out_label = out_label; f.out_checksum = reinterpret_cast<unsigned long long *>(out_checksum);
  f.out_id = out_id; f.out_dist = out_dist;
  return launch_finalize(f, st(stream));
}

int dmlp_synth_attrs(double *dst, int64_t rows, int a, int64_t row_base, double lo, double hi, uint64_t seed,
                     uint64_t stream_id, void *stream) {
  return launch_synth_attrs(dst, rows, a, row_base, lo, hi, seed, stream_id, st(stream));
}
int dmlp_synth_labels(int *dst, int64_t rows, int64_t row_base, int num_labels, uint64_t seed, void *stream) {
  return launch_synth_labels(dst, rows, row_base, num_labels, seed, st(stream));
}
int dmlp_synth_k(int *dst, int64_t rows, int64_t row_base, int min_k, int max_k, uint64_t seed, void *stream) {
  return launch_synth_k(dst, rows, row_base, min_k, max_k, seed, st(stream));
}

// ---- whole local pipeline ------------------------------------------------------
namespace {
struct LocalLayout {
  size_t stats, q_stats, xh, qh, xn, xe, qn, qe, cand, rerank, rerank_bytes, total;
  int a_pad, chunks;
};
LocalLayout local_layout(int q, int n, int a, int num_sms, int kmax) {
  LocalLayout L;
  L.a_pad = pad64(a > 0 ? a : 1);
  TopkPlan p = make_topk_plan(q, n, L.a_pad, 32, num_sms, kmax);
  L.chunks = p.lists;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += up256(bytes); return o; };
  L.stats = take(DMLP_STATS_BYTES);
  L.q_stats = take(DMLP_STATS_BYTES);
  L.xh = take(size_t(n) * L.a_pad * 2);
  L.qh = take(size_t(q) * L.a_pad * 2);
  L.xn = take(size_t(n) * 4);
  L.xe = take(size_t(n) * 4);
  L.qn = take(size_t(q) * 4);
  L.qe = take(size_t(q) * 4);
  L.cand = take(size_t(q) * p.lists * 32 * 8);
  L.rerank_bytes = rerank_workspace_bytes(q, kmax);
  L.rerank = take(L.rerank_bytes);
  L.total = off;
  return L;
}
}  // namespace

size_t dmlp_knn_local_workspace_bytes(int q, int n, int a, int num_sms, int kmax) {
  if (num_sms <= 0) num_sms = 148;
  return local_layout(q, n, a, num_sms, kmax).total;
}

int dmlp_knn_local(const double *q_attrs, const int *k, int q, const double *x_attrs, const int *x_labels, int n,
                   int a, int id_base, int kmax, const void *range, void *workspace, size_t workspace_bytes,
                   double *out_dist, int *out_id, int *out_label, int *flags, int *num_flagged, void *stream) {
  DMLP_REQUIRE(q >= 0 && n >= 0 && a >= 0 && kmax >= 1, "knn_local: negative size or kmax < 1");
  int sms = 0;
  if (int rc = current_sm_count(&sms)) return rc;
  LocalLayout L = local_layout(q, n, a, sms, kmax);
  DMLP_REQUIRE(workspace_bytes >= L.total, "knn_local: workspace too small");
  DMLP_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "knn_local: workspace must be 256-byte aligned");
  unsigned char *w = static_cast<unsigned char *>(workspace);
  cudaStream_t s = st(stream);
  if (num_flagged) DMLP_CUDA_OK(cudaMemsetAsync(num_flagged, 0, sizeof(int), s));
  if (q <= 0) return 0;
  if (n <= 0) {
    // nothing to filter or re-rank (and the scratch may hold another call's candidates)
    const long long slots = static_cast<long long>(q) * kmax;
    long long blocks = (slots + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    fill_empty_lists_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(out_dist, out_id, out_label, flags, slots, q);
    DMLP_CUDA_OK(cudaGetLastError());
    return 0;
  }
  PackStats *stats = reinterpret_cast<PackStats *>(w + L.stats);
  PackStats *q_stats = reinterpret_cast<PackStats *>(w + L.q_stats);
  if (int rc = launch_stats_reset(stats, s)) return rc;
  if (int rc = launch_stats_reset(q_stats, s)) return rc;
  const PackStats *rng = static_cast<const PackStats *>(range);
  if (rng == nullptr) {
    if (int rc = launch_stats(x_attrs, static_cast<long long>(n) * a, stats, s)) return rc;
    if (int rc = launch_stats(q_attrs, static_cast<long long>(q) * a, stats, s)) return rc;
    rng = stats;
  }
  __half *xh = reinterpret_cast<__half *>(w + L.xh), *qh = reinterpret_cast<__half *>(w + L.qh);
  float *xn = reinterpret_cast<float *>(w + L.xn), *xe = reinterpret_cast<float *>(w + L.xe);
  float *qn = reinterpret_cast<float *>(w + L.qn), *qe = reinterpret_cast<float *>(w + L.qe);
  unsigned long long *cand = reinterpret_cast<unsigned long long *>(w + L.cand);
  // the row maxima live in the stats blocks (range fields of q_stats stay unused)
  if (int rc = launch_pack(x_attrs, n, a, L.a_pad, rng, xh, xn, xe, stats, s)) return rc;
  if (int rc = launch_pack(q_attrs, q, a, L.a_pad, rng, qh, qn, qe, q_stats, s)) return rc;
  TopkPlan p = make_topk_plan(q, n, L.a_pad, 32, sms, kmax);
  if (int rc = launch_dist_topk(qh, xh, xn, p, cand, nullptr, s)) return rc;
  RerankArgs r;
  r.cand = cand; r.chunks = p.lists; r.kp = 32;
  r.q_attrs = q_attrs; r.x_attrs = x_attrs; r.x_labels = x_labels; r.k = k;
  r.q_norm2 = qn; r.q_errn = qe; r.range = rng; r.x_maxima = stats;
  r.q = q; r.n = n; r.a = a; r.a_pad = L.a_pad; r.id_base = id_base; r.kmax = kmax;
  r.out_dist = out_dist; r.out_id = out_id; r.out_label = out_label; r.flags = flags; r.num_flagged = num_flagged;
  r.workspace = w + L.rerank; r.workspace_bytes = L.rerank_bytes;
  return launch_rerank(r, s);
}

}  // extern "C"
