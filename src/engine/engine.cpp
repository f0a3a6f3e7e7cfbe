// Engine::KNN / KNNFlat: backend selection and the reporting loop.
//
// Reference flow (engine.cpp:20-350): broadcast sizes, build a process grid,
// scatter/broadcast shards, brute-force locally, gather k-lists, vote, report.
// Here the flow is: flat views -> backend (GPU pipeline or CPU strategy) ->
// ResultSet -> one reportResult() per query.
#include "engine.h"

#include <cstdlib>
#include <cstring>
#include <stdexcept>

#include "dmlp/gpu.h"

namespace dmlp {

static int env_int(const char *name, int dflt) {
  const char *v = std::getenv(name);
  return (v && *v) ? std::atoi(v) : dflt;
}

EngineConfig EngineConfig::from_env() {
  EngineConfig c;
  if (const char *b = std::getenv("DMLP_BACKEND")) {
    std::string s(b);
    if (s == "cpu") c.backend = Backend::kCpu;
    else if (s == "gpu" || s == "gpu_fused") c.backend = Backend::kGpuFused;
    else if (s == "gpu_exact") c.backend = Backend::kGpuExact;
    else c.backend = Backend::kAuto;
  }
  if (const char *s = std::getenv("DMLP_STRATEGY")) {
    Strategy st;
    if (!parse_strategy(s, &st)) throw std::invalid_argument(std::string("unknown DMLP_STRATEGY: ") + s);
    c.cpu.strategy = st;
  }
  c.cpu.ranks = env_int("DMLP_RANKS", 1);
  c.cpu.threads = env_int("DMLP_THREADS", 0);
  if (const char *g = std::getenv("DMLP_GRID")) {
    int r = 0, cc = 0;
    if (std::sscanf(g, "%dx%d", &r, &cc) == 2 && r > 0 && cc > 0) {
      c.grid_rows = r; c.grid_cols = cc;
      c.cpu.rows = r; c.cpu.cols = cc;
      if (!std::getenv("DMLP_RANKS")) c.cpu.ranks = r * cc;
    } else {
      throw std::invalid_argument(std::string("DMLP_GRID must look like RxC, got: ") + g);
    }
  }
  c.ngpu = env_int("DMLP_NGPU", 0);
  c.trace = env_int("DMLP_TRACE", 0) != 0;
  return c;
}

}  // namespace dmlp

using namespace dmlp;

void Engine::KNNFlat(const DatasetView &data, const QueryView &queries, ResultSet *out) {
  ResultSet local;
  ResultSet *rs = out ? out : &local;

  Backend be = config.backend;
  int ndev = (be == Backend::kCpu) ? 0 : gpu::device_count();
  if (be == Backend::kAuto) be = ndev > 0 ? Backend::kGpuFused : Backend::kCpu;
  if ((be == Backend::kGpuFused || be == Backend::kGpuExact) && ndev == 0)
    throw std::runtime_error("GPU backend requested but no CUDA device is visible");

  if (be == Backend::kCpu) {
    knn_cpu(data, queries, config.cpu, rs);
    last_backend = std::string("cpu:") + strategy_name(config.cpu.strategy);
  } else {
    gpu::GpuOptions go;
    go.ngpu = config.ngpu > 0 ? (config.ngpu < ndev ? config.ngpu : ndev) : ndev;
    go.rows = config.grid_rows;
    go.cols = config.grid_cols;
    go.exact_only = (be == Backend::kGpuExact);
    go.trace = config.trace;
    go.have_range = config.have_range;
    go.range_lo = config.range_lo;
    go.range_hi = config.range_hi;
    go.wire_i32 = data.attrs_i32 != nullptr && queries.attrs_i32 != nullptr;
    gpu::knn_gpu(data, queries, go, rs, &last_backend);
  }

  if (!out) {
    // Exactly one report per query (the reference prints each P times on P
    // ranks; that defect is not reproduced).
    Query shell;
    std::vector<std::pair<double, int>> pairs;
    for (int64_t i = 0; i < queries.q; ++i) {
      shell.id = queries.id_base + static_cast<int>(i);
      shell.k = queries.k[i];
      pairs.clear();
      for (int64_t j = rs->offset[i]; j < rs->offset[i + 1]; ++j)
        pairs.emplace_back(rs->dist[j], rs->ids[j]);
      reportResult(shell, pairs, rs->label[i]);
    }
  }
}

void Engine::Prepare(const Params &p, int kmax) {
  if (config.backend == Backend::kCpu) return;
  const int ndev = gpu::device_count();
  if (ndev == 0) return;
  gpu::GpuOptions go;
  go.ngpu = config.ngpu > 0 ? (config.ngpu < ndev ? config.ngpu : ndev) : ndev;
  go.rows = config.grid_rows;
  go.cols = config.grid_cols;
  go.exact_only = (config.backend == Backend::kGpuExact);
  go.wire_i32 = config.wire_i32;
  gpu::prepare(p.num_data, p.num_queries, p.num_attrs, kmax, go);
}

void Engine::KNN(Params &p, std::vector<DataPoint> &dataset, std::vector<Query> &queries) {
  // AoS -> SoA at the API boundary (the reference does the same repack right
  // before its scatters, engine.cpp:79-96 and 154-184).
  const int a = p.num_attrs;
  const int64_t n = static_cast<int64_t>(dataset.size());
  const int64_t q = static_cast<int64_t>(queries.size());
  const bool want_pinned = config.backend != Backend::kCpu && gpu::device_count() > 0;
  Problem flat(want_pinned ? gpu::pinned_allocator() : malloc_allocator());
  flat.resize(n, q, a);
  bool ids_are_rows = true;
  for (int64_t i = 0; i < n; ++i) {
    const DataPoint &d = dataset[size_t(i)];
    if (d.id != i) ids_are_rows = false;
    flat.data_labels[i] = d.label;
    size_t have = d.attrs.size() < size_t(a) ? d.attrs.size() : size_t(a);
    std::memcpy(flat.data_attrs + i * a, d.attrs.data(), have * sizeof(double));
    for (size_t j = have; j < size_t(a); ++j) flat.data_attrs[i * a + j] = 0.0;
  }
  for (int64_t i = 0; i < q; ++i) {
    const Query &qq = queries[size_t(i)];
    flat.query_k[i] = qq.k;
    size_t have = qq.attrs.size() < size_t(a) ? qq.attrs.size() : size_t(a);
    std::memcpy(flat.query_attrs + i * a, qq.attrs.data(), have * sizeof(double));
    for (size_t j = have; j < size_t(a); ++j) flat.query_attrs[i * a + j] = 0.0;
  }

  ResultSet rs;
  KNNFlat(flat.dataset(), flat.queries(), &rs);

  // Map row indices back to the caller's ids and report.
  std::vector<std::pair<double, int>> pairs;
  for (int64_t i = 0; i < q; ++i) {
    pairs.clear();
    std::vector<int> ids;
    for (int64_t j = rs.offset[i]; j < rs.offset[i + 1]; ++j) {
      int row = rs.ids[j];
      int id = (row >= 0 && !ids_are_rows) ? dataset[size_t(row)].id : row;
      pairs.emplace_back(rs.dist[j], id);
    }
    reportResult(queries[size_t(i)], pairs, rs.label[i]);
  }
}
