// CPU backends (oracle + multi-"rank" strategies).  See include/dmlp/cpu.h.
//
// Build note: this file must be compiled with -ffp-contract=off so that the
// distance loop keeps the reference's mul-then-add rounding.
#include "dmlp/cpu.h"

#include <algorithm>
#include <atomic>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <thread>

namespace dmlp {

// ---------------------------------------------------------------------------
// small shared helpers
// ---------------------------------------------------------------------------

int vote_label(const Neighbor *nb, int64_t k) {
  std::vector<int> labels;
  labels.reserve(static_cast<size_t>(k));
  for (int64_t i = 0; i < k; ++i)
    if (nb[i].id >= 0) labels.push_back(nb[i].label);
  if (labels.empty()) return -1;
  std::sort(labels.begin(), labels.end());
  int best = labels[0], best_count = 0;
  size_t i = 0;
  while (i < labels.size()) {
    size_t j = i;
    while (j < labels.size() && labels[j] == labels[i]) ++j;
    int count = static_cast<int>(j - i);
    if (count >= best_count) {  // ascending scan + ">=" => larger label wins ties
      best_count = count;
      best = labels[i];
    }
    i = j;
  }
  return best;
}

void choose_grid(int ranks, int *rows, int *cols) {
  if (ranks < 1) ranks = 1;
  int c = 1;
  for (int d = 1; d * d <= ranks; ++d)
    if (ranks % d == 0) c = d;
  *cols = c;
  *rows = ranks / c;
}

bool parse_strategy(const std::string &name, Strategy *out) {
  static const struct { const char *n; Strategy s; } table[] = {
      {"serial", Strategy::kSerial},   {"grid", Strategy::kGrid2D},
      {"gather", Strategy::kShardGather}, {"tree", Strategy::kShardTreeReduce},
      {"farm", Strategy::kQueryFarm},  {"kdtree", Strategy::kKdTree}};
  for (auto &e : table)
    if (name == e.n) { *out = e.s; return true; }
  return false;
}

const char *strategy_name(Strategy s) {
  switch (s) {
    case Strategy::kSerial: return "serial";
    case Strategy::kGrid2D: return "grid";
    case Strategy::kShardGather: return "gather";
    case Strategy::kShardTreeReduce: return "tree";
    case Strategy::kQueryFarm: return "farm";
    case Strategy::kKdTree: return "kdtree";
  }
  return "?";
}

double exact_sqdist(const double *a, const double *b, int dim) {
  double acc = 0.0;
  for (int i = 0; i < dim; ++i) {
    double d = a[i] - b[i];
    double sq = d * d;
    acc = acc + sq;
  }
  return acc;
}

namespace {

// Four independent accumulation chains (four different points) so the oracle is
// not bound by the latency of one add chain; each chain is still sequential.
inline void exact_sqdist_x4(const double *q, const double *p0, const double *p1,
                            const double *p2, const double *p3, int dim, double out[4]) {
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  for (int i = 0; i < dim; ++i) {
    double qi = q[i];
    double d0 = qi - p0[i], d1 = qi - p1[i], d2 = qi - p2[i], d3 = qi - p3[i];
    double m0 = d0 * d0, m1 = d1 * d1, m2 = d2 * d2, m3 = d3 * d3;
    s0 = s0 + m0; s1 = s1 + m1; s2 = s2 + m2; s3 = s3 + m3;
  }
  out[0] = s0; out[1] = s1; out[2] = s2; out[3] = s3;
}

// Max-heap on `closer`: the root is the farthest of the kept candidates.
struct BoundedHeap {
  Neighbor *h;
  int cap;
  int size = 0;
  BoundedHeap(Neighbor *storage, int k) : h(storage), cap(k) {}
  inline void offer(const Neighbor &c) {
    if (cap == 0) return;
    if (size < cap) {
      h[size++] = c;
      std::push_heap(h, h + size, closer);
    } else if (closer(c, h[0])) {
      std::pop_heap(h, h + size, closer);
      h[size - 1] = c;
      std::push_heap(h, h + size, closer);
    }
  }
  inline bool full() const { return size == cap; }
  inline const Neighbor &worst() const { return h[0]; }
  // Sort ascending and pad with sentinels up to capacity.
  void finish() {
    std::sort_heap(h, h + size, closer);
    for (int i = size; i < cap; ++i) h[i] = Neighbor{kSentinelDist, kSentinelId, -1};
  }
};

void parallel_for(int64_t n, int threads, const std::function<void(int64_t, int64_t, int)> &fn) {
  if (threads <= 0) threads = static_cast<int>(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  if (n < threads) threads = static_cast<int>(std::max<int64_t>(n, 1));
  if (threads == 1) { fn(0, n, 0); return; }
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    int64_t b = n * t / threads, e = n * (t + 1) / threads;
    pool.emplace_back([=, &fn] { fn(b, e, t); });
  }
  for (auto &th : pool) th.join();
}

DatasetView data_slice(const DatasetView &d, int64_t begin, int64_t count) {
  DatasetView s = d;
  s.n = count;
  s.attrs = d.attrs + begin * d.a;
  s.labels = d.labels + begin;
  s.id_base = d.id_base + static_cast<int>(begin);
  return s;
}

int clamp_k(int k) { return k < 0 ? 0 : k; }

// CSR offsets of the per-query k's.
std::vector<int64_t> k_offsets(const QueryView &qv) {
  std::vector<int64_t> off(static_cast<size_t>(qv.q) + 1, 0);
  for (int64_t i = 0; i < qv.q; ++i) off[i + 1] = off[i] + clamp_k(qv.k[i]);
  return off;
}

}  // namespace

void local_topk(const DatasetView &data, const double *query, int k, Neighbor *out) {
  k = clamp_k(k);
  BoundedHeap heap(out, k);
  int64_t r = 0;
  const int a = data.a;
  for (; r + 4 <= data.n; r += 4) {
    double d[4];
    const double *p = data.attrs + r * a;
    exact_sqdist_x4(query, p, p + a, p + 2 * a, p + 3 * a, a, d);
    for (int j = 0; j < 4; ++j)
      heap.offer(Neighbor{d[j], data.id_base + static_cast<int>(r + j), data.labels[r + j]});
  }
  for (; r < data.n; ++r)
    heap.offer(Neighbor{exact_sqdist(query, data.attrs + r * a, a),
                        data.id_base + static_cast<int>(r), data.labels[r]});
  heap.finish();
}

namespace {

constexpr int kPointGroup = 8;   // points per transposed group (one AVX-512 register of doubles)
constexpr int kQueryBlock = 64;  // queries that share one pass over the data

// Distances of queries to eight points held attribute-major (`pt[i * 8 + j]` = attribute i of
// point j).  The eight points are the SIMD dimension, written with GCC vector types so that the
// compiler cannot pick another one; every lane runs exact_sqdist's sequence (the file is
// compiled with -ffp-contract=off: no fused multiply-add in any clone).  Four queries at a
// time give four independent add chains per loaded point vector.
typedef double v8d __attribute__((vector_size(64), aligned(8)));

__attribute__((target_clones("avx512f", "avx2", "default")))
void sqdist_group8(const double *q, const double *pt, int dim, double *out) {
  v8d s = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < dim; ++i) {
    const v8d p = *reinterpret_cast<const v8d *>(pt + static_cast<size_t>(i) * kPointGroup);
    const v8d d = q[i] - p;
    const v8d m = d * d;
    s = s + m;
  }
  *reinterpret_cast<v8d *>(out) = s;
}

__attribute__((target_clones("avx512f", "avx2", "default")))
void sqdist_group8_x4(const double *q0, const double *q1, const double *q2, const double *q3, const double *pt,
                      int dim, double *out) {
  v8d s0 = {0, 0, 0, 0, 0, 0, 0, 0}, s1 = s0, s2 = s0, s3 = s0;
  for (int i = 0; i < dim; ++i) {
    const v8d p = *reinterpret_cast<const v8d *>(pt + static_cast<size_t>(i) * kPointGroup);
    const v8d d0 = q0[i] - p, d1 = q1[i] - p, d2 = q2[i] - p, d3 = q3[i] - p;
    const v8d m0 = d0 * d0, m1 = d1 * d1, m2 = d2 * d2, m3 = d3 * d3;
    s0 = s0 + m0;
    s1 = s1 + m1;
    s2 = s2 + m2;
    s3 = s3 + m3;
  }
  v8d *o = reinterpret_cast<v8d *>(out);
  o[0] = s0; o[1] = s1; o[2] = s2; o[3] = s3;
}

}  // namespace

void local_topk_many(const DatasetView &data, const double *queries, const int *k, int nq, Neighbor *const *outs) {
  const int a = data.a;
  std::vector<double> pt(static_cast<size_t>(a > 0 ? a : 1) * kPointGroup);
  std::vector<BoundedHeap> heaps;
  for (int b = 0; b < nq; b += kQueryBlock) {
    const int nb = std::min(kQueryBlock, nq - b);
    heaps.clear();
    for (int i = 0; i < nb; ++i) heaps.emplace_back(outs[b + i], clamp_k(k[b + i]));
    int64_t r = 0;
    for (; r + kPointGroup <= data.n; r += kPointGroup) {
      const double *rows = data.attrs + r * a;
      for (int j = 0; j < kPointGroup; ++j)
        for (int i = 0; i < a; ++i) pt[static_cast<size_t>(i) * kPointGroup + j] = rows[static_cast<size_t>(j) * a + i];
      int i = 0;
      double d[4 * kPointGroup];
      const double *qrow = queries + static_cast<size_t>(b) * a;
      for (; i + 4 <= nb; i += 4) {
        sqdist_group8_x4(qrow + static_cast<size_t>(i) * a, qrow + static_cast<size_t>(i + 1) * a,
                         qrow + static_cast<size_t>(i + 2) * a, qrow + static_cast<size_t>(i + 3) * a, pt.data(), a, d);
        for (int u = 0; u < 4; ++u) {
          BoundedHeap &h = heaps[static_cast<size_t>(i + u)];
          for (int j = 0; j < kPointGroup; ++j) {
            const double dj = d[u * kPointGroup + j];
            if (h.cap == 0 || (h.full() && dj > h.worst().dist)) continue;  // (an equal distance can still win on the id)
            h.offer(Neighbor{dj, data.id_base + static_cast<int>(r + j), data.labels[r + j]});
          }
        }
      }
      for (; i < nb; ++i) {
        sqdist_group8(qrow + static_cast<size_t>(i) * a, pt.data(), a, d);
        for (int j = 0; j < kPointGroup; ++j)
          heaps[static_cast<size_t>(i)].offer(
              Neighbor{d[j], data.id_base + static_cast<int>(r + j), data.labels[r + j]});
      }
    }
    for (; r < data.n; ++r)
      for (int i = 0; i < nb; ++i)
        heaps[static_cast<size_t>(i)].offer(Neighbor{exact_sqdist(queries + static_cast<size_t>(b + i) * a,
                                                                  data.attrs + r * a, a),
                                                     data.id_base + static_cast<int>(r), data.labels[r]});
    for (auto &h : heaps) h.finish();
  }
}

void merge_topk(const Neighbor *x, const Neighbor *y, int k, Neighbor *out) {
  int i = 0, j = 0;
  for (int o = 0; o < k; ++o) {
    if (closer(y[j], x[i])) out[o] = y[j++];
    else out[o] = x[i++];
  }
}

void finalize_query(const Neighbor *nb, int k, ResultSet *out) {
  if (out->offset.empty()) out->offset.push_back(0);
  int label = vote_label(nb, k);
  size_t base = out->ids.size();
  for (int i = 0; i < k; ++i) {
    out->dist.push_back(nb[i].dist);
    out->ids.push_back(nb[i].id);
  }
  out->offset.push_back(static_cast<int64_t>(out->ids.size()));
  out->label.push_back(label);
  out->checksum.push_back(result_checksum(label, out->ids.data() + base, k));
}

namespace {

// Final lists for all queries (CSR by k) -> ResultSet, in query order.
void finalize_all(const QueryView &qv, const std::vector<int64_t> &off,
                  const std::vector<Neighbor> &lists, ResultSet *out) {
  out->clear();
  out->offset.push_back(0);
  for (int64_t i = 0; i < qv.q; ++i)
    finalize_query(lists.data() + off[i], static_cast<int>(off[i + 1] - off[i]), out);
}

namespace {
// Lists of queries [b, e) into `lists` (CSR by `off`), through the blocked kernel.
void topk_range(const DatasetView &data, const QueryView &qv, const std::vector<int64_t> &off, int64_t b, int64_t e,
                Neighbor *lists) {
  if (e <= b) return;
  std::vector<Neighbor *> outs(static_cast<size_t>(e - b));
  for (int64_t i = b; i < e; ++i) outs[static_cast<size_t>(i - b)] = lists + off[i];
  local_topk_many(data, qv.attrs + b * qv.a, qv.k + b, static_cast<int>(e - b), outs.data());
}
}  // namespace

// ---- strategy: serial / oracle -------------------------------------------
void run_serial(const DatasetView &data, const QueryView &qv, int threads, ResultSet *out) {
  auto off = k_offsets(qv);
  std::vector<Neighbor> lists(static_cast<size_t>(off.back()));
  parallel_for(qv.q, threads, [&](int64_t b, int64_t e, int) { topk_range(data, qv, off, b, e, lists.data()); });
  finalize_all(qv, off, lists, out);
}

// ---- strategy: R x C grid (covers gather and tree as C == 1 special cases) --
//
// Rank (r, c) owns data shard r and query shard c.  Phase 1: every rank writes
// its per-query sorted, sentinel-padded k-lists into its own mailbox.  Phase 2
// (the "reduction" along the data axis, reference: engine.cpp:273-309) runs per
// grid column, either as gather-then-select at row 0 or as a pairwise tree of
// `merge_topk` operators.
enum class Reduce { kGatherSelect, kTree };

void run_grid(const DatasetView &data, const QueryView &qv, int rows, int cols, Reduce how,
              ResultSet *out) {
  BlockPartition dp{data.n, rows}, qp{qv.q, cols};
  auto off = k_offsets(qv);
  // mailbox[r] holds lists for ALL queries (each column fills its own slice).
  std::vector<std::vector<Neighbor>> mailbox(static_cast<size_t>(rows));
  for (auto &m : mailbox) m.resize(static_cast<size_t>(off.back()));

  std::vector<std::thread> ranks;
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c)
      ranks.emplace_back([&, r, c] {
        DatasetView shard = data_slice(data, dp.begin(r), dp.count(r));
        int64_t qb = qp.begin(c), qe = qb + qp.count(c);
        topk_range(shard, qv, off, qb, qe, mailbox[r].data());
      });
  for (auto &t : ranks) t.join();

  std::vector<Neighbor> final_lists(static_cast<size_t>(off.back()));
  std::vector<std::thread> owners;
  for (int c = 0; c < cols; ++c)
    owners.emplace_back([&, c] {
      int64_t qb = qp.begin(c), qe = qb + qp.count(c);
      std::vector<Neighbor> scratch, tmp;
      for (int64_t i = qb; i < qe; ++i) {
        int k = static_cast<int>(off[i + 1] - off[i]);
        Neighbor *dst = final_lists.data() + off[i];
        if (how == Reduce::kGatherSelect) {
          scratch.clear();
          for (int r = 0; r < rows; ++r)
            scratch.insert(scratch.end(), mailbox[r].data() + off[i], mailbox[r].data() + off[i] + k);
          std::partial_sort(scratch.begin(), scratch.begin() + k, scratch.end(), closer);
          std::copy(scratch.begin(), scratch.begin() + k, dst);
        } else {
          // pairwise tree: stride doubling, result accumulates in row 0's slot
          tmp.resize(static_cast<size_t>(k));
          for (int stride = 1; stride < rows; stride *= 2)
            for (int r = 0; r + stride < rows; r += 2 * stride) {
              Neighbor *x = mailbox[r].data() + off[i];
              const Neighbor *y = mailbox[r + stride].data() + off[i];
              merge_topk(x, y, k, tmp.data());
              std::copy(tmp.begin(), tmp.end(), x);
            }
          std::copy(mailbox[0].data() + off[i], mailbox[0].data() + off[i] + k, dst);
        }
      }
    });
  for (auto &t : owners) t.join();
  finalize_all(qv, off, final_lists, out);
}

// ---- strategy: dynamic query farm ------------------------------------------
// Every worker sees the whole dataset and pulls the next query index from a
// shared counter, the shared-memory analogue of bench_4's master handing out
// one query at a time over Send/Recv.
void run_farm(const DatasetView &data, const QueryView &qv, int workers, ResultSet *out) {
  auto off = k_offsets(qv);
  std::vector<Neighbor> lists(static_cast<size_t>(off.back()));
  std::atomic<int64_t> next{0};
  if (workers < 1) workers = 1;
  // requests small enough to balance (about four per worker), large enough to share data passes
  const int64_t grain = std::max<int64_t>(1, std::min<int64_t>(kQueryBlock, qv.q / (4 * int64_t(workers))));
  std::vector<std::thread> pool;
  for (int w = 0; w < workers; ++w)
    pool.emplace_back([&] {
      for (;;) {  // a block of queries per request: they share one pass over the data
        int64_t i = next.fetch_add(grain, std::memory_order_relaxed);
        if (i >= qv.q) break;
        topk_range(data, qv, off, i, std::min<int64_t>(i + grain, qv.q), lists.data());
      }
    });
  for (auto &t : pool) t.join();
  finalize_all(qv, off, lists, out);
}

// ---- strategy: exact kd-tree -------------------------------------------------
// Median-split tree over row indices.  Distances are still evaluated with
// exact_sqdist so results are bit-identical to brute force; pruning compares the
// squared plane offset against the current worst kept distance and only prunes
// on strict ">" because an equal-distance point with a larger id still wins.
struct KdTree {
  struct Node { int axis; int row; int left; int right; };
  const DatasetView &d;
  std::vector<Node> nodes;
  std::vector<int> order;
  explicit KdTree(const DatasetView &data) : d(data) {
    order.resize(static_cast<size_t>(d.n));
    for (int64_t i = 0; i < d.n; ++i) order[i] = static_cast<int>(i);
    nodes.reserve(order.size());
    if (d.n > 0) build(0, static_cast<int>(d.n), 0);
  }
  int build(int lo, int hi, int depth) {
    if (lo >= hi) return -1;
    int axis = d.a > 0 ? depth % d.a : 0;
    int mid = lo + (hi - lo) / 2;
    std::nth_element(order.begin() + lo, order.begin() + mid, order.begin() + hi,
                     [&](int x, int y) {
                       return d.attrs[int64_t(x) * d.a + axis] < d.attrs[int64_t(y) * d.a + axis];
                     });
    int me = static_cast<int>(nodes.size());
    nodes.push_back(Node{axis, order[mid], -1, -1});
    int l = build(lo, mid, depth + 1);
    int r = build(mid + 1, hi, depth + 1);
    nodes[me].left = l;
    nodes[me].right = r;
    return me;
  }
  void search(int ni, const double *q, BoundedHeap &heap) const {
    if (ni < 0) return;
    const Node &nd = nodes[ni];
    const double *p = d.attrs + int64_t(nd.row) * d.a;
    heap.offer(Neighbor{exact_sqdist(q, p, d.a), d.id_base + nd.row, d.labels[nd.row]});
    double diff = q[nd.axis] - p[nd.axis];
    int near = diff < 0 ? nd.left : nd.right;
    int far = diff < 0 ? nd.right : nd.left;
    search(near, q, heap);
    double plane = diff * diff;
    if (heap.cap > 0 && (!heap.full() || !(plane > heap.worst().dist))) search(far, q, heap);  // (k == 0: nothing to find)
  }
};

void run_kdtree(const DatasetView &data, const QueryView &qv, ResultSet *out) {
  auto off = k_offsets(qv);
  std::vector<Neighbor> lists(static_cast<size_t>(off.back()));
  KdTree tree(data);
  for (int64_t i = 0; i < qv.q; ++i) {
    BoundedHeap heap(lists.data() + off[i], static_cast<int>(off[i + 1] - off[i]));
    if (heap.cap == 0) { heap.finish(); continue; }
    if (data.a > 0) tree.search(data.n > 0 ? 0 : -1, qv.attrs + i * qv.a, heap);
    else for (int64_t r = 0; r < data.n; ++r)
      heap.offer(Neighbor{0.0, data.id_base + static_cast<int>(r), data.labels[r]});
    heap.finish();
  }
  finalize_all(qv, off, lists, out);
}

}  // namespace

void knn_cpu(const DatasetView &data, const QueryView &queries, const CpuOptions &opt,
             ResultSet *out) {
  if (data.a != queries.a && queries.q > 0 && data.n > 0)
    throw std::invalid_argument("dataset and queries disagree on num_attrs");
  int ranks = opt.ranks < 1 ? 1 : opt.ranks;
  switch (opt.strategy) {
    case Strategy::kSerial:
      run_serial(data, queries, opt.threads, out);
      return;
    case Strategy::kGrid2D: {
      int rows = opt.rows, cols = opt.cols;
      if (rows <= 0 || cols <= 0) choose_grid(ranks, &rows, &cols);
      run_grid(data, queries, rows, cols, Reduce::kGatherSelect, out);
      return;
    }
    case Strategy::kShardGather:
      run_grid(data, queries, ranks, 1, Reduce::kGatherSelect, out);
      return;
    case Strategy::kShardTreeReduce:
      run_grid(data, queries, ranks, 1, Reduce::kTree, out);
      return;
    case Strategy::kQueryFarm:
      run_farm(data, queries, ranks, out);
      return;
    case Strategy::kKdTree:
      run_kdtree(data, queries, out);
      return;
  }
}

}  // namespace dmlp
