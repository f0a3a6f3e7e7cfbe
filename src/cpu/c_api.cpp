// C ABI over the CPU backends and the loader.  See include/dmlp_cpu.h.
#include "dmlp_cpu.h"

#include <cstring>
#include <cstdio>
#include <exception>
#include <stdexcept>
#include <thread>
#include <vector>
#include <string>

#include "dmlp/cpu.h"
#include "dmlp/io.h"

namespace {
thread_local std::string g_err;
template <class F>
int guarded(F &&f) {
  try { f(); return 0; }
  catch (const std::exception &e) { g_err = e.what(); return -1; }
  catch (...) { g_err = "unknown error"; return -1; }
}
// A Problem that borrows caller memory: allocator hands out the caller's
// arrays in the order Problem::resize asks for them and never frees.
thread_local void *g_borrow[4];
thread_local int g_borrow_next;
dmlp::HostAllocator borrowing_allocator() {
  dmlp::HostAllocator a;
  a.alloc = [](size_t) -> void * { return g_borrow[g_borrow_next++ & 3]; };
  a.release = [](void *) {};
  return a;
}
}  // namespace

extern "C" {

const char *dmlp_cpu_last_error(void) { return g_err.c_str(); }

int dmlp_cpu_knn(const double *data_attrs, const int *data_labels, int64_t n,
                 const double *query_attrs, const int *query_k, int64_t q, int a,
                 const char *strategy, int ranks, int rows, int cols, int threads,
                 int *out_label, uint64_t *out_checksum, int *out_ids, double *out_dist) {
  return guarded([&] {
    dmlp::DatasetView d; d.n = n; d.a = a; d.attrs = data_attrs; d.labels = data_labels;
    dmlp::QueryView qv; qv.q = q; qv.a = a; qv.attrs = query_attrs; qv.k = query_k;
    dmlp::CpuOptions opt;
    if (!dmlp::parse_strategy(strategy ? strategy : "serial", &opt.strategy))
      throw std::invalid_argument(std::string("unknown strategy: ") + strategy);
    opt.ranks = ranks; opt.rows = rows; opt.cols = cols; opt.threads = threads;
    dmlp::ResultSet rs;
    dmlp::knn_cpu(d, qv, opt, &rs);
    if (out_label) std::memcpy(out_label, rs.label.data(), sizeof(int) * size_t(q));
    if (out_checksum) std::memcpy(out_checksum, rs.checksum.data(), sizeof(uint64_t) * size_t(q));
    if (out_ids) std::memcpy(out_ids, rs.ids.data(), sizeof(int) * rs.ids.size());
    if (out_dist) std::memcpy(out_dist, rs.dist.data(), sizeof(double) * rs.dist.size());
  });
}

double dmlp_exact_sqdist(const double *a, const double *b, int dim) {
  return dmlp::exact_sqdist(a, b, dim);
}

uint64_t dmlp_result_checksum(int label, const int *ids, int64_t k) {
  return dmlp::result_checksum(label, ids, k);
}

void dmlp_merge_topk(const double *xd, const int *xi, const double *yd, const int *yi, int k,
                     double *od, int *oi) {
  const size_t kk = static_cast<size_t>(k);
  std::vector<dmlp::Neighbor> x(kk), y(kk), o(kk);
  for (int i = 0; i < k; ++i) { x[i] = {xd[i], xi[i], 0}; y[i] = {yd[i], yi[i], 0}; }
  dmlp::merge_topk(x.data(), y.data(), k, o.data());
  for (int i = 0; i < k; ++i) { od[i] = o[i].dist; oi[i] = o[i].id; }
}

int dmlp_cpu_local_lists(const double *data_attrs, const int *data_labels, int64_t n, int id_base,
                         const double *query_attrs, const int *query_k, int64_t q, int a, int kmax, int threads,
                         double *out_dist, int *out_id, int *out_label) {
  return guarded([&] {
    dmlp::DatasetView d; d.n = n; d.a = a; d.attrs = data_attrs; d.labels = data_labels; d.id_base = id_base;
    if (threads <= 0) threads = int(std::thread::hardware_concurrency());
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
      pool.emplace_back([&, t] {
        const int64_t qb = q * t / threads, qe = q * (t + 1) / threads;
        const size_t stride = static_cast<size_t>(kmax > 0 ? kmax : 1);
        std::vector<dmlp::Neighbor> tmp(stride * static_cast<size_t>(qe - qb));
        std::vector<dmlp::Neighbor *> outs(static_cast<size_t>(qe - qb));
        std::vector<int> kk(static_cast<size_t>(qe - qb));
        for (int64_t i = qb; i < qe; ++i) {
          kk[size_t(i - qb)] = query_k[i] < 0 ? 0 : (query_k[i] > kmax ? kmax : query_k[i]);
          outs[size_t(i - qb)] = tmp.data() + stride * size_t(i - qb);
        }
        dmlp::local_topk_many(d, query_attrs + qb * a, kk.data(), int(qe - qb), outs.data());
        for (int64_t i = qb; i < qe; ++i) {
          const int k = kk[size_t(i - qb)];
          const dmlp::Neighbor *row = outs[size_t(i - qb)];
          for (int j = 0; j < kmax; ++j) {
            const bool real = j < k;
            out_dist[i * kmax + j] = real ? row[j].dist : dmlp::kSentinelDist;
            out_id[i * kmax + j] = real ? row[j].id : dmlp::kSentinelId;
            out_label[i * kmax + j] = real ? row[j].label : -1;
          }
        }
      });
    for (auto &th : pool) th.join();
  });
}

int dmlp_cpu_finalize_lists(const double *dist, const int *id, const int *label, int sources, int64_t q, int kmax,
                            const int *query_k, int *out_label, uint64_t *out_checksum) {
  return guarded([&] {
    std::vector<dmlp::Neighbor> acc, next, cur;
    for (int64_t i = 0; i < q; ++i) {
      int k = query_k[i] < 0 ? 0 : (query_k[i] > kmax ? kmax : query_k[i]);
      const size_t kk = static_cast<size_t>(k);
      acc.assign(kk, dmlp::Neighbor{dmlp::kSentinelDist, dmlp::kSentinelId, -1});
      next.resize(kk); cur.resize(kk);
      for (int s = 0; s < sources; ++s) {
        const int64_t base = (int64_t(s) * q + i) * kmax;
        for (int j = 0; j < k; ++j) cur[size_t(j)] = dmlp::Neighbor{dist[base + j], id[base + j], label[base + j]};
        dmlp::merge_topk(acc.data(), cur.data(), k, next.data());   // the reduction operator, applied source by source
        acc.swap(next);
      }
      dmlp::ResultSet rs;
      dmlp::finalize_query(acc.data(), k, &rs);
      out_label[i] = rs.label[0];
      out_checksum[i] = rs.checksum[0];
    }
  });
}

int dmlp_parse_header(const char *text, size_t len, int64_t *n, int64_t *q, int *a) {
  return guarded([&] {
    long long nn = 0, qq = 0; int aa = 0;
    std::string head(text, len < 256 ? len : 256);
    if (std::sscanf(head.c_str(), "%lld %lld %d", &nn, &qq, &aa) != 3)
      throw std::runtime_error("malformed header line");
    *n = nn; *q = qq; *a = aa;
  });
}

int dmlp_parse_text(const char *text, size_t len, int threads, double *data_attrs, int *data_labels,
                    double *query_attrs, int *query_k) {
  return guarded([&] {
    g_borrow[0] = data_attrs; g_borrow[1] = data_labels;
    g_borrow[2] = query_attrs; g_borrow[3] = query_k;
    g_borrow_next = 0;
    dmlp::Problem p(borrowing_allocator());
    dmlp::parse_problem(text, len, threads, &p);
  });
}

int dmlp_encode_decimal6_host(const double *values, int64_t count, int threads, int32_t *out, int *ok) {
  return guarded([&] { *ok = dmlp::encode_decimal6_values(values, count, threads, out) ? 1 : 0; });
}

int dmlp_decimal6_bits_host(const int32_t *values, int64_t count, int *bits) {
  return guarded([&] { *bits = dmlp::decimal6_bits(values, count); });
}

int dmlp_pack_bits_host(const int32_t *src, int64_t rows, int a, int bits, int words, uint32_t *dst) {
  return guarded([&] { dmlp::pack_bits_rows(src, rows, a, bits, words, dst); });
}

int dmlp_unpack_bits_host(const uint32_t *src, int64_t rows, int a, int bits, int words, int32_t *dst) {
  return guarded([&] { dmlp::unpack_bits_rows(src, rows, a, bits, words, dst); });
}

int dmlp_parse_update(const char *text, size_t len, int *id, double *attrs, int capacity, int *count) {
  return guarded([&] {
    std::vector<double> v;
    if (!dmlp::parse_update_line(text, len, id, &v)) throw std::runtime_error("Line is empty");
    *count = static_cast<int>(v.size());
    for (int i = 0; i < *count && i < capacity; ++i) attrs[i] = v[size_t(i)];
  });
}

int dmlp_synth_fill(int64_t n, int64_t q, int a, double lo, double hi, int min_k, int max_k,
                    int num_labels, uint64_t seed, double *data_attrs, int *data_labels,
                    double *query_attrs, int *query_k) {
  return guarded([&] {
    g_borrow[0] = data_attrs; g_borrow[1] = data_labels;
    g_borrow[2] = query_attrs; g_borrow[3] = query_k;
    g_borrow_next = 0;
    dmlp::Problem p(borrowing_allocator());
    dmlp::SynthSpec s;
    s.n = n; s.q = q; s.a = a; s.lo = lo; s.hi = hi; s.min_k = min_k; s.max_k = max_k;
    s.num_labels = num_labels; s.seed = seed;
    dmlp::synth_problem(s, &p);
  });
}

void dmlp_choose_grid(int ranks, int *rows, int *cols) { dmlp::choose_grid(ranks, rows, cols); }
int64_t dmlp_block_begin(int64_t total, int parts, int p) {
  return dmlp::BlockPartition{total, parts}.begin(p);
}
int64_t dmlp_block_count(int64_t total, int parts, int p) {
  return dmlp::BlockPartition{total, parts}.count(p);
}

}  // extern "C"
