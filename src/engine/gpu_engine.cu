// GPU backend of Engine::KNNFlat: one host process drives an R x C grid of B200s with the
// same fused kernels the one-process-per-GPU Python engine uses.
//
// Reference flow being replaced (engine.cpp:40-309): MPI_Dims_create grid, Scatterv + Bcast of
// data shards along rows and query shards along columns, local brute force, MPI_Gather of
// k-lists to row 0, re-select.  Here:
//
//   * the runtime is persistent: streams, peer access between every pair of devices
//     (cudaDeviceEnablePeerAccess) and ONE grow-only device arena per GPU are set up once
//     (Engine::Prepare / the first call); a call carves its buffers out of the arena, so the
//     timed region contains no cudaMalloc / cudaFree;
//   * every device gets only ITS share of the inputs from (pinned) host memory -- 1/C of its
//     data shard on a 1 x C grid, 1/R of its column's queries when R > 1 -- and the rest arrives
//     over NVLink inside the filter GEMM: the communication warps of dist_topk pull the peers'
//     packed rows through plain UVA peer pointers (K4, the reference's row / column MPI_Bcast);
//   * the exact k-lists go straight from the re-rank kernel into the owner's receive buffer
//     (K5, the reference's MPI_Gather + re-select); "packed", "done reading" and "lists landed"
//     are flag words written by the last block of the producing kernel and awaited inside the
//     consuming kernel -- no cudaMemcpyPeerAsync, no signal / wait launches, no host sync
//     between the first copy and the final read-back;
//   * queries the filter cannot certify are repaired on the device (dmlp_repair_flagged).
// Block partitions give the remainder to shard 0 like the reference (engine.cpp:62-63,
// 136-137) and non-square grids are handled correctly.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "dmlp/cpu.h"
#include "dmlp/gpu.h"
#include "multicast.h"
#include "dmlp_knn.h"

namespace dmlp {
namespace gpu {

namespace {

void check(cudaError_t e, const char *what) {
  if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}
void check_rc(int rc, const char *what) {
  if (rc != 0) throw std::runtime_error(std::string(what) + ": " + dmlp_last_error());
}
#define CK(expr) check((expr), #expr)

constexpr int kFusedMaxK = 96;   // keep in sync with ops/cuda.py FUSED_MAX_K
constexpr int kRepairCap = 64;   // flagged queries a call repairs on the device (ops/cuda.py REPAIR_CAP)
constexpr int kMaxDev = 8;

inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }
inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }
inline int pad64(int a) { return (std::max(a, 1) + 63) / 64 * 64; }

// ---- persistent per-process runtime ---------------------------------------------------------
struct Device {
  int id = 0;
  cudaStream_t stream = nullptr;
  unsigned char *arena = nullptr;
  size_t arena_bytes = 0;
  int sms = 148;
};

class Runtime {
 public:
  static Runtime &get() {
    static Runtime rt;
    return rt;
  }
  // streams + peer access for devices [0, n)
  void ensure_devices(int n) {
    std::lock_guard<std::mutex> hold(lock_);
    while (static_cast<int>(devs_.size()) < n) {
      Device d;
      d.id = static_cast<int>(devs_.size());
      CK(cudaSetDevice(d.id));
      CK(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
      CK(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, d.id));
      devs_.push_back(d);
    }
    if (n > peers_) {
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
          if (i == j || (i < peers_ && j < peers_)) continue;
          int can = 0;
          CK(cudaDeviceCanAccessPeer(&can, i, j));
          if (!can) throw std::runtime_error("GPUs " + std::to_string(i) + " and " + std::to_string(j) +
                                             " cannot access each other's memory (no NVLink / P2P)");
          CK(cudaSetDevice(i));
          cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
          if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
          else check(e, "cudaDeviceEnablePeerAccess");
        }
      peers_ = n;
    }
  }
  // the arena of device d holds at least `bytes` (grows outside the steady state; contents are
  // per call, except the flag words at its head, which are zeroed whenever it is (re)allocated)
  unsigned char *reserve(int d, size_t bytes) {
    Device &dev = devs_[static_cast<size_t>(d)];
    if (dev.arena_bytes < bytes) {
      CK(cudaSetDevice(dev.id));
      CK(cudaDeviceSynchronize());
      if (dev.arena) CK(cudaFree(dev.arena));
      dev.arena = nullptr;
      dev.arena_bytes = 0;
      const size_t want = bytes + bytes / 8;
      CK(cudaMalloc(reinterpret_cast<void **>(&dev.arena), want));
      dev.arena_bytes = want;
      CK(cudaMemset(dev.arena, 0, 4096));
      ++generation_;
    }
    return dev.arena;
  }
  Device &dev(int d) { return devs_[static_cast<size_t>(d)]; }
  // NVSwitch multicast pages for the flags of an all-device group (multicast.h); created once per
  // device count, kept for the life of the process; ->ok is false where multicast is not on offer
  McPages *mc_pages(int ndev) {
    std::lock_guard<std::mutex> hold(lock_);
    auto it = mc_.find(ndev);
    if (it == mc_.end()) {
      it = mc_.emplace(ndev, McPages::create(ndev, 4096)).first;
      if (!it->second.ok && std::getenv("DMLP_TRACE"))
        std::fprintf(stderr, "[dmlp] NVLS flags off for %d GPUs: %s\n", ndev, it->second.why.c_str());
    }
    return &it->second;
  }
  unsigned next_epoch() { return ++epoch_; }
  // flag words live in the arenas: a re-allocation zeroes them, so epochs restart as well
  unsigned long long generation() const { return generation_; }
  void reset_epoch() { epoch_ = 0; }

 private:
  std::mutex lock_;
  std::vector<Device> devs_;
  std::map<int, McPages> mc_;
  int peers_ = 0;
  unsigned epoch_ = 0;
  unsigned long long generation_ = 0;
};

// ---- buffer layout of one call (identical on every device, so a peer's buffer is the peer's
// arena base plus the same offset: symmetric memory without IPC) -----------------------------
struct Shape {
  int R = 1, C = 1, a = 0, a_pad = 64, kmax = 1;
  int64_t n_max = 0;    // rows of the largest data shard
  int64_t q_max = 0;    // queries of the largest column
  int64_t qs = 0;       // slice stride: queries per owner inside a column (padded to 256 when R > 1)
  int64_t qp = 0;       // R * qs
  bool fused = true;
  bool wire32 = false;  // inputs cross PCIe as decimal6 integers (views' attrs_i32) and are expanded on the device
};

struct Layout {
  // head: flag arrays and counters (zeroed when the arena is created, monotonic epochs afterwards)
  size_t pack_flags = 0, list_flags = 128, done_flags = 256, counters = 384, status = 512, nflag = 516;
  size_t stats = 768, q_stats = 1024;
  size_t x64, labels, xh, xn, xe, q64, k, qh, qn, qe, cand, rr_ws, fix_ws, od, oi, ol, flags, frows, final_label,
      final_sum, fid, fdist, xw = 0, qw = 0, tile_count, rb_count, total;
  int chunks = 2;
  size_t rr_bytes = 0, fix_bytes = 0;
};

Layout make_layout(const Shape &s, int sms) {
  Layout L;
  size_t off = 4096;
  auto take = [&](size_t bytes) { size_t o = off; off += up256(bytes ? bytes : 1); return o; };
  const size_t n = static_cast<size_t>(s.n_max), qp = static_cast<size_t>(s.qp), km = static_cast<size_t>(s.kmax);
  L.x64 = take(n * s.a * 8);
  L.labels = take(n * 4);
  L.q64 = take(qp * s.a * 8);
  L.k = take(qp * 4);
  if (s.fused) {
    L.xh = take(n * s.a_pad * 2);
    L.xn = take(n * 4);
    L.xe = take(n * 4);
    L.qh = take(qp * s.a_pad * 2);
    L.qn = take(qp * 4);
    L.qe = take(qp * 4);
    int plan[8];
    check_rc(dmlp_topk_plan(static_cast<int>(qp), static_cast<int>(std::max<int64_t>(s.n_max, 1)), s.a_pad, sms, s.kmax,
                            plan), "topk_plan");
    L.chunks = plan[3];
    L.cand = take(qp * L.chunks * 32 * 8);
    L.rr_bytes = dmlp_rerank_workspace_bytes(static_cast<int>(qp), s.kmax);
    L.rr_ws = take(L.rr_bytes);
    L.fix_bytes = dmlp_exact_workspace_bytes(static_cast<int>(s.n_max), kRepairCap, s.kmax);
  } else {
    L.xh = L.xn = L.xe = L.qh = L.qn = L.qe = L.cand = L.rr_ws = 0;
    L.fix_bytes = dmlp_exact_workspace_bytes(static_cast<int>(s.n_max), static_cast<int>(qp), s.kmax);
  }
  L.fix_ws = take(L.fix_bytes);
  // lists: local [qp][kmax] with one data shard, receive buffers [R][qs][kmax] otherwise
  const size_t slots = (s.R > 1 ? static_cast<size_t>(s.R) * s.qs : qp) * km;
  L.od = take(slots * 8);
  L.oi = take(slots * 4);
  L.ol = take(slots * 4);
  L.flags = take(qp * 4);
  L.frows = take(kRepairCap * 4);
  L.final_label = take(static_cast<size_t>(s.qs) * 4);
  L.final_sum = take(static_cast<size_t>(s.qs) * 8);
  L.fid = take(static_cast<size_t>(s.qs) * km * 4);
  L.fdist = take(static_cast<size_t>(s.qs) * km * 8);
  if (s.wire32) {   // staging for the int32 form of the attributes (same element offsets as x64 / q64)
    L.xw = take(n * s.a * 4);
    L.qw = take(qp * s.a * 4);
  }
  L.tile_count = take((n / 256 + 2) * 4);
  L.rb_count = take((qp / 128 + 2) * 4);
  L.total = off;
  return L;
}

}  // namespace

int device_count() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

HostAllocator pinned_allocator() {
  if (device_count() == 0) return malloc_allocator();
  HostAllocator a;
  a.alloc = [](size_t bytes) -> void * {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
  };
  a.release = [](void *p) { cudaFreeHost(p); };
  return a;
}

namespace {

struct Plan {
  Shape shape;
  int ndev = 1;
  std::vector<Layout> layout;   // one per device (same offsets; SM counts could differ in theory)
  size_t bytes = 0;
  bool mc = false;              // "tell every peer" goes through NVSwitch multicast pages (one store, replicated)
  std::string key;
};

// Decide grid, path and buffer sizes for a problem.
Plan make_plan(int64_t N, int64_t Q, int a, int kmax, const GpuOptions &opt) {
  Plan p;
  const int have = std::max(1, std::min(std::min(opt.ngpu, device_count()), kMaxDev));
  int R = opt.rows, C = opt.cols;
  if (R <= 0 || C <= 0 || R * C > have) { R = have; C = 1; }
  Shape &s = p.shape;
  s.R = R; s.C = C; s.a = a; s.a_pad = pad64(a); s.kmax = std::max(kmax, 1);
  s.n_max = BlockPartition{N, R}.count(0);
  s.q_max = BlockPartition{Q, C}.count(0);
  // several devices gather operands inside the filter kernel: k <= 96 there, k <= 1024 on one device
  s.fused = !opt.exact_only && N >= R && Q > 0 &&   // (no empty data shard on the fused path)
            dmlp_fused_supported(kmax, N / R) >= (R * C > 1 ? 2 : 1) &&
            double(N) * double(Q) * double(std::max(a, 1)) >= double(1 << 22);
  // slices of a column: one per data shard; padded to whole 256-row tiles for the fused gather
  const int64_t per = BlockPartition{s.q_max, R}.count(0);
  s.qs = R > 1 ? round_up(std::max<int64_t>(per, 1), 256) : std::max<int64_t>(s.q_max, 1);
  s.qp = s.qs * R;
  s.wire32 = opt.wire_i32;
  // the fused data-row gather needs the CTA-pair kernel (>= 256 queries per column)
  p.ndev = R * C;
  Runtime &rt = Runtime::get();
  rt.ensure_devices(p.ndev);
  for (int d = 0; d < p.ndev; ++d) {
    p.layout.push_back(make_layout(s, rt.dev(d).sms));
    p.bytes = std::max(p.bytes, p.layout.back().total);
  }
  // the signalling groups (a column when R > 1, a row when C > 1) are all devices when the other
  // dimension is 1: then one multicast object serves them (DMLP_NVLS=0 keeps per-peer stores)
  const char *nvls = std::getenv("DMLP_NVLS");
  if (p.ndev > 1 && (R == 1 || C == 1) && !(nvls && std::string(nvls) == "0")) p.mc = rt.mc_pages(p.ndev)->ok;
  char buf[176];
  std::snprintf(buf, sizeof buf, "%dx%d a=%d k=%d n=%lld q=%lld N=%lld Q=%lld f=%d w=%d m=%d", R, C, a, s.kmax,
                (long long)s.n_max, (long long)s.qp, (long long)N, (long long)Q, int(s.fused), int(s.wire32), int(p.mc));
  p.key = buf;
  return p;
}

std::string g_last_key;
unsigned long long g_last_generation = ~0ULL;

// Arenas in place, flag words and counters valid for this shape.
void prepare_plan(const Plan &p) {
  Runtime &rt = Runtime::get();
  for (int d = 0; d < p.ndev; ++d) rt.reserve(d, p.bytes);
  if (p.key != g_last_key || rt.generation() != g_last_generation) {
    // a new shape (or new arenas): the monotonic flag / counter protocol starts over
    for (int d = 0; d < p.ndev; ++d) {
      CK(cudaSetDevice(rt.dev(d).id));
      CK(cudaDeviceSynchronize());
    }
    for (int d = 0; d < p.ndev; ++d) {
      const Layout &L = p.layout[size_t(d)];
      CK(cudaSetDevice(rt.dev(d).id));
      CK(cudaMemset(rt.dev(d).arena, 0, 4096));
      CK(cudaMemset(rt.dev(d).arena + L.tile_count, 0, L.rb_count - L.tile_count));
      CK(cudaMemset(rt.dev(d).arena + L.rb_count, 0, L.total - L.rb_count));
      if (p.mc) CK(cudaMemset(rt.mc_pages(p.ndev)->local[d], 0, 4096));
      CK(cudaDeviceSynchronize());
    }
    rt.reset_epoch();
    g_last_key = p.key;
    g_last_generation = rt.generation();
  }
}

}  // namespace

void prepare(int64_t num_data, int64_t num_queries, int num_attrs, int kmax, const GpuOptions &opt) {
  if (device_count() == 0) return;
  Plan p = make_plan(num_data, num_queries, num_attrs, kmax, opt);
  prepare_plan(p);
}

namespace {

struct Owner {   // what one device reports back
  int64_t qb = 0, q = 0;   // global query range it owns
};

template <class T> T *at(unsigned char *base, size_t off) { return reinterpret_cast<T *>(base + off); }

void fill_signal(dmlp_signal_args *sg, const std::vector<unsigned char *> &bases, size_t off_flags, int world, int rank,
                 unsigned value, unsigned char *own_base, size_t off_counters, int counter_slot) {
  std::memset(sg, 0, sizeof *sg);
  for (int i = 0; i < world; ++i) sg->peer_flags[i] = bases[size_t(i)] + off_flags;
  sg->world = world; sg->rank = rank; sg->value = value;
  sg->done_counter = at<uint32_t>(own_base, off_counters) + counter_slot;
}

void fill_wait(dmlp_wait_args *w, unsigned char *own_base, size_t off_flags, int world, unsigned value, size_t off_status) {
  w->flags = at<uint32_t>(own_base, off_flags);
  w->world = world; w->value = value; w->timeout_seconds = 20.0;
  w->status = at<int>(own_base, off_status);
}

}  // namespace

void knn_gpu(const DatasetView &data, const QueryView &queries, const GpuOptions &opt, ResultSet *out,
             std::string *path) {
  const int a = data.a;
  const int64_t N = data.n, Q = queries.q;
  int kmax = 1;
  for (int64_t i = 0; i < Q; ++i) kmax = std::max(kmax, queries.k[i]);

  out->clear();
  out->offset.assign(size_t(Q) + 1, 0);
  for (int64_t i = 0; i < Q; ++i) out->offset[i + 1] = out->offset[i] + std::max(queries.k[i], 0);
  out->dist.assign(size_t(out->offset[Q]), 0.0);
  out->ids.assign(size_t(out->offset[Q]), -1);
  out->label.assign(size_t(Q), -1);
  out->checksum.assign(size_t(Q), 0);

  Plan plan = make_plan(N, Q, a, kmax, opt);
  const Shape &S = plan.shape;
  if (S.wire32 && (data.attrs_i32 == nullptr || queries.attrs_i32 == nullptr))
    throw std::runtime_error("gpu: wire_i32 needs attrs_i32 in both views");
  const int R = S.R, C = S.C;
  if (path) {
    char buf[96];
    std::snprintf(buf, sizeof buf, "gpu:%s R=%d C=%d%s", S.fused ? "fused" : "exact", R, C, plan.mc ? " nvls" : "");
    *path = buf;
  }
  if (Q == 0) return;
  auto t0 = std::chrono::steady_clock::now();
  prepare_plan(plan);   // no-op when Engine::Prepare (or an earlier call of this shape) already did it
  Runtime &rt = Runtime::get();
  const unsigned epoch = rt.next_epoch();

  // value range of everything (the loader's when it supplied one): common centring / scaling
  double lo = opt.range_lo, hi = opt.range_hi;
  if (S.fused && !(opt.have_range)) {
    lo = 1e300; hi = -1e300;
    const int64_t nd = N * a, nq = Q * a;
    for (int64_t i = 0; i < nd; ++i) { const double v = data.attrs[i]; lo = v < lo ? v : lo; hi = v > hi ? v : hi; }
    for (int64_t i = 0; i < nq; ++i) { const double v = queries.attrs[i]; lo = v < lo ? v : lo; hi = v > hi ? v : hi; }
  }

  BlockPartition dp{N, R}, qp{Q, C};
  std::vector<unsigned char *> base(size_t(R) * C);
  for (int d = 0; d < R * C; ++d) base[size_t(d)] = rt.dev(d).arena;
  auto dev_of = [&](int r, int c) { return r * C + c; };
  // multicast flag pages: list flags at 0, "done reading" flags at 128 (word = rank in the group = device)
  const McPages *mcp = plan.mc ? rt.mc_pages(plan.ndev) : nullptr;
  constexpr size_t kMcList = 0, kMcDone = 128;
  // rows_fused: one data shard whose parts are spread over the row and gathered inside the GEMM
  const char *rf_env = std::getenv("DMLP_ROWS_FUSED");   // 0: every device loads the whole shard from the host
  const bool rows_fused = S.fused && R == 1 && C > 1 && qp.count(C - 1) >= 256 && N >= C &&
                          !(rf_env && std::atoi(rf_env) == 0);
  const int km = S.kmax;

  std::vector<std::vector<int>> h_nflag(size_t(R) * C, std::vector<int>(1, 0));
  std::vector<std::vector<int>> h_ids(size_t(R) * C);
  std::vector<std::vector<double>> h_dist(size_t(R) * C);
  std::vector<Owner> owned(size_t(R) * C);
  std::vector<int> h_status(size_t(R) * C, 0);
  struct Readback { bool folded = false; int64_t qb = 0, q = 0; };
  std::vector<Readback> readback(size_t(R) * C);
  // ---- enqueue everything, device by device; nothing here waits for the GPU ----------------
  for (int r = 0; r < R; ++r)
    for (int c = 0; c < C; ++c) {
      const int d = dev_of(r, c);
      Device &dv = rt.dev(d);
      const Layout &L = plan.layout[size_t(d)];
      unsigned char *B = base[size_t(d)];
      cudaStream_t st = dv.stream;
      CK(cudaSetDevice(dv.id));
      const int64_t n = dp.count(r), nb = dp.begin(r);     // data shard of this row
      const int64_t qc = qp.count(c), qcb = qp.begin(c);   // queries of this column
      BlockPartition sl{qc, R};                            // their owners along the column
      const int64_t own_q = R > 1 ? sl.count(r) : qc, own_b = R > 1 ? sl.begin(r) : 0;
      double *x64 = at<double>(B, L.x64), *q64 = at<double>(B, L.q64);
      int *xl = at<int>(B, L.labels), *qk = at<int>(B, L.k);
      int *nflag = at<int>(B, L.nflag), *flags = at<int>(B, L.flags), *frows = at<int>(B, L.frows);
      int *f_label = at<int>(B, L.final_label);
      uint64_t *f_sum = at<uint64_t>(B, L.final_sum);
      std::vector<unsigned char *> col, row;
      col.resize(static_cast<size_t>(R));
      row.resize(static_cast<size_t>(C));
      for (int rr = 0; rr < R; ++rr) col[size_t(rr)] = base[size_t(dev_of(rr, c))];
      for (int cc = 0; cc < C; ++cc) row[size_t(cc)] = base[size_t(dev_of(r, cc))];

      // -- host -> device: only this device's share
      BlockPartition parts{n, C};   // rows_fused: part c of the shard
      const int64_t pb = rows_fused ? parts.begin(c) : 0, pn = rows_fused ? parts.count(c) : n;
      // attributes: doubles, or (wire32) the loader's decimal6 integers -- half the bytes -- expanded
      // on the device into the very doubles the parser produced (dec6_to_double, common.cuh)
      int *xw = S.wire32 ? at<int>(B, L.xw) : nullptr, *qw = S.wire32 ? at<int>(B, L.qw) : nullptr;
      auto h2d_attrs = [&](double *dst, int *stage, const double *src64, const int32_t *src32, int64_t rows) {
        const size_t count = size_t(rows) * size_t(a);
        if (count == 0) return;
        if (S.wire32) {
          CK(cudaMemcpyAsync(stage, src32, sizeof(int32_t) * count, cudaMemcpyHostToDevice, st));
          check_rc(dmlp_expand_decimal6(stage, dst, int64_t(count), st), "expand_decimal6");
        } else {
          CK(cudaMemcpyAsync(dst, src64, sizeof(double) * count, cudaMemcpyHostToDevice, st));
        }
      };
      if (pn > 0) {
        h2d_attrs(x64 + pb * a, xw ? xw + pb * a : nullptr, data.attrs + (nb + pb) * a,
                  S.wire32 ? data.attrs_i32 + (nb + pb) * a : nullptr, pn);
        CK(cudaMemcpyAsync(xl + pb, data.labels + nb + pb, sizeof(int) * size_t(pn), cudaMemcpyHostToDevice, st));
      }
      if (R > 1) {
        // padded column index space: slice rr at [rr * qs, ...).  Fused: only the own slice travels
        // (the GEMM gathers the others); exact path: every slice comes from the host.
        CK(cudaMemsetAsync(qk, 0, sizeof(int) * size_t(S.qp), st));
        for (int rr = 0; rr < R; ++rr) {
          if (S.fused && rr != r) continue;
          const int64_t cnt = sl.count(rr), bg = sl.begin(rr);
          if (cnt == 0) continue;
          h2d_attrs(q64 + rr * S.qs * a, qw ? qw + rr * S.qs * a : nullptr, queries.attrs + (qcb + bg) * a,
                    S.wire32 ? queries.attrs_i32 + (qcb + bg) * a : nullptr, cnt);
          CK(cudaMemcpyAsync(qk + rr * S.qs, queries.k + qcb + bg, sizeof(int) * size_t(cnt), cudaMemcpyHostToDevice, st));
        }
      } else if (qc > 0) {
        h2d_attrs(q64, qw, queries.attrs + qcb * a, S.wire32 ? queries.attrs_i32 + qcb * a : nullptr, qc);
        CK(cudaMemcpyAsync(qk, queries.k + qcb, sizeof(int) * size_t(qc), cudaMemcpyHostToDevice, st));
      }
      CK(cudaMemsetAsync(nflag, 0, sizeof(int), st));

      dmlp_scatter_args sc;
      std::memset(&sc, 0, sizeof sc);
      sc.world = R; sc.rank = r; sc.slice_rows = int(S.qs);
      for (int rr = 0; rr < R; ++rr) sc.peer_base[rr] = col[size_t(rr)];
      sc.off_dist = L.od; sc.off_id = L.oi; sc.off_label = L.ol;
      dmlp_signal_args sg_lists, sg_pack, sg_done;
      dmlp_wait_args w_lists, w_done;
      fill_signal(&sg_lists, col, L.list_flags, R, r, epoch, B, L.counters, 0);
      fill_wait(&w_lists, B, L.list_flags, R, epoch, L.status);
      if (mcp != nullptr && R > 1) {   // (C == 1: the column is every device)
        for (int rr = 0; rr < R; ++rr) sg_lists.peer_flags[rr] = mcp->local[dev_of(rr, c)] + kMcList;
        sg_lists.mc_flags = reinterpret_cast<uint32_t *>(mcp->mc + kMcList);
        w_lists.flags = reinterpret_cast<const uint32_t *>(mcp->local[d] + kMcList);
      }

      if (!S.fused) {
        // ---- direct fp64 path: whole column's queries against this shard
        dmlp_exact_args e;
        std::memset(&e, 0, sizeof e);
        e.q_attrs = q64; e.x_attrs = x64; e.x_labels = xl; e.k = qk; e.count = int(R > 1 ? S.qp : qc);
        e.n = int(n); e.a = a; e.id_base = data.id_base + int(nb); e.kmax = km;
        e.out_dist = at<double>(B, L.od); e.out_id = at<int>(B, L.oi); e.out_label = at<int>(B, L.ol);
        e.workspace = B + L.fix_ws; e.workspace_bytes = L.fix_bytes;
        if (R > 1) { e.scatter = &sc; e.signal = &sg_lists; }
        check_rc(dmlp_exact_ex(&e, st), "exact");
      } else {
        void *stats = B + L.stats, *q_stats = B + L.q_stats;
        check_rc(dmlp_stats_reset(stats, st), "stats_reset");
        check_rc(dmlp_stats_reset(q_stats, st), "stats_reset");
        check_rc(dmlp_stats_set(stats, lo, hi, st), "stats_set");
        void *xh = B + L.xh, *qh = B + L.qh;
        float *xn = at<float>(B, L.xn), *xe = at<float>(B, L.xe), *qn = at<float>(B, L.qn), *qe = at<float>(B, L.qe);
        uint64_t *cand = at<uint64_t>(B, L.cand);
        dmlp_rerank_args rr_args;
        std::memset(&rr_args, 0, sizeof rr_args);
        if (R > 1) {
          // K4 + K5 along the column
          check_rc(dmlp_pack(x64, n, a, S.a_pad, stats, xh, xn, xe, stats, st), "pack(x)");
          fill_signal(&sg_pack, col, L.pack_flags, R, r, epoch, B, L.counters, 1);
          check_rc(dmlp_pack_ex(q64 + r * S.qs * a, own_q, a, S.a_pad, stats,
                                static_cast<unsigned char *>(qh) + size_t(r) * S.qs * S.a_pad * 2, qn + r * S.qs,
                                qe + r * S.qs, q_stats, nullptr, &sg_pack, st), "pack(q)");
          dmlp_gather_args g;
          std::memset(&g, 0, sizeof g);
          g.world = R; g.rank = r; g.slice_rows = int(S.qs); g.a = a; g.epoch = epoch;
          for (int rr = 0; rr < R; ++rr) g.peer_base[rr] = col[size_t(rr)];
          g.off_qh = L.qh; g.off_q64 = L.q64; g.off_qn = L.qn; g.off_qe = L.qe; g.off_k = L.k;
          g.off_pack_flags = L.pack_flags;
          g.rb_count = at<uint32_t>(B, L.rb_count); g.status = at<int>(B, L.status); g.timeout_seconds = 20.0;
          check_rc(dmlp_dist_topk_gather(xh, xn, int(n), S.a_pad, dv.sms, km, cand, &g, st), "dist_topk_gather");
          rr_args.scatter = &sc;
        } else if (rows_fused) {
          // fused row all-gather of the data shard (1 x C)
          fill_signal(&sg_pack, row, L.pack_flags, C, c, epoch, B, L.counters, 1);
          fill_signal(&sg_done, row, L.done_flags, C, c, epoch, B, L.counters, 2);
          fill_wait(&w_done, B, L.done_flags, C, epoch - 1, L.status);
          if (mcp != nullptr) {   // (R == 1: the row is every device)
            for (int cc = 0; cc < C; ++cc) sg_done.peer_flags[cc] = mcp->local[dev_of(r, cc)] + kMcDone;
            sg_done.mc_flags = reinterpret_cast<uint32_t *>(mcp->mc + kMcDone);
            w_done.flags = reinterpret_cast<const uint32_t *>(mcp->local[d] + kMcDone);
          }
          check_rc(dmlp_pack_ex(x64 + pb * a, pn, a, S.a_pad, stats, static_cast<unsigned char *>(xh) + size_t(pb) * S.a_pad * 2,
                                xn + pb, xe + pb, nullptr, &w_done, &sg_pack, st), "pack(x part)");
          check_rc(dmlp_pack(q64, qc, a, S.a_pad, stats, qh, qn, qe, q_stats, st), "pack(q)");
          dmlp_data_gather_args g;
          std::memset(&g, 0, sizeof g);
          g.world = C; g.rank = c; g.a = a; g.epoch = epoch;
          for (int cc = 0; cc < C; ++cc) { g.peer_base[cc] = row[size_t(cc)]; g.part_begin[cc] = int(parts.begin(cc)); }
          g.part_begin[C] = int(n);
          g.off_xh = L.xh; g.off_xn = L.xn; g.off_xe = L.xe; g.off_x64 = L.x64; g.off_labels = L.labels;
          g.off_pack_flags = L.pack_flags;
          g.tile_count = at<uint32_t>(B, L.tile_count); g.status = at<int>(B, L.status); g.timeout_seconds = 20.0;
          g.done_signal = &sg_done;
          check_rc(dmlp_dist_topk_rows(qh, int(qc), int(n), S.a_pad, dv.sms, km, cand, &g, st), "dist_topk_rows");
          check_rc(dmlp_maxima_from_norms(xn, xe, n, stats, st), "maxima_from_norms");
        } else {
          check_rc(dmlp_pack(x64, n, a, S.a_pad, stats, xh, xn, xe, stats, st), "pack(x)");
          check_rc(dmlp_pack(q64, qc, a, S.a_pad, stats, qh, qn, qe, q_stats, st), "pack(q)");
          check_rc(dmlp_dist_topk(qh, xh, xn, int(qc), int(n), S.a_pad, dv.sms, km, cand, st), "dist_topk");
        }
        const int q_rows = int(R > 1 ? S.qp : qc);
        int plan_now[8];
        check_rc(dmlp_topk_plan(q_rows, int(std::max<int64_t>(n, 1)), S.a_pad, dv.sms, km, plan_now), "topk_plan");
        rr_args.cand = cand; rr_args.chunks = plan_now[3]; rr_args.kp = 32;
        rr_args.q_attrs = q64; rr_args.x_attrs = x64; rr_args.x_labels = xl; rr_args.k = qk;
        rr_args.q_norm2 = qn; rr_args.q_errn = qe; rr_args.range = stats; rr_args.x_maxima = stats;
        rr_args.q = q_rows; rr_args.n = int(n); rr_args.a = a; rr_args.a_pad = S.a_pad;
        rr_args.id_base = data.id_base + int(nb); rr_args.kmax = km;
        rr_args.flags = flags; rr_args.num_flagged = nflag; rr_args.flagged_rows = frows; rr_args.flagged_cap = kRepairCap;
        rr_args.workspace = B + L.rr_ws; rr_args.workspace_bytes = L.rr_bytes;
        dmlp_repair_args fx;
        std::memset(&fx, 0, sizeof fx);
        fx.q_attrs = q64; fx.x_attrs = x64; fx.x_labels = xl; fx.k = qk; fx.flagged_rows = frows; fx.num_flagged = nflag;
        fx.cap = kRepairCap; fx.q = q_rows; fx.n = int(n); fx.a = a; fx.id_base = data.id_base + int(nb); fx.kmax = km;
        fx.workspace = B + L.fix_ws; fx.workspace_bytes = L.fix_bytes;
        if (R > 1) {
          fx.scatter = &sc; fx.signal = &sg_lists;
        } else {
          rr_args.out_dist = fx.out_dist = at<double>(B, L.od);
          rr_args.out_id = fx.out_id = at<int>(B, L.oi);
          rr_args.out_label = fx.out_label = at<int>(B, L.ol);
          rr_args.final_label = fx.final_label = f_label;
          rr_args.final_checksum = fx.final_checksum = f_sum;
        }
        if (n > 0) {
          check_rc(dmlp_rerank_ex(&rr_args, st), "rerank");
          check_rc(dmlp_repair_flagged(&fx, st), "repair_flagged");
        } else {   // an empty shard contributes padding only (and still has to announce it)
          dmlp_exact_args e;
          std::memset(&e, 0, sizeof e);
          e.q_attrs = q64; e.x_attrs = x64; e.x_labels = xl; e.k = qk; e.count = q_rows; e.n = 0; e.a = a;
          e.id_base = data.id_base + int(nb); e.kmax = km;
          e.out_dist = at<double>(B, L.od); e.out_id = at<int>(B, L.oi); e.out_label = at<int>(B, L.ol);
          e.workspace = B + L.fix_ws; e.workspace_bytes = L.fix_bytes;
          if (R > 1) { e.scatter = &sc; e.signal = &sg_lists; }
          check_rc(dmlp_exact_ex(&e, st), "exact(empty shard)");
        }
      }
      // ---- owner side: merge the column's lists (waiting for them inside the kernel), vote, checksum
      if (own_q > 0) {
        const bool folded = S.fused && R == 1 && n > 0;
        if (!folded) {
          dmlp_finalize_args f;
          std::memset(&f, 0, sizeof f);
          f.dist = at<double>(B, L.od); f.id = at<int>(B, L.oi); f.label = at<int>(B, L.ol);
          f.source_stride = int64_t(S.qs) * km; f.sources = R; f.q = int(R > 1 ? S.qs : qc); f.kmax = km;
          f.k = qk + (R > 1 ? r * S.qs : 0);
          f.out_label = f_label; f.out_checksum = f_sum;
          f.out_id = at<int>(B, L.fid); f.out_dist = at<double>(B, L.fdist);
          if (R > 1) f.wait = &w_lists;
          check_rc(dmlp_finalize_ex(&f, st), "finalize");
        }
        Readback &rb = readback[size_t(d)];
        rb.folded = folded;
        rb.qb = qcb + own_b;
        rb.q = own_q;
      }
    }
  // ---- read-backs, once every device has all of its work (a copy into pageable host memory
  // blocks the host until the device gets there, and an owner's finalize waits for lists that
  // the other devices' kernels -- all enqueued above -- produce)
  for (int d = 0; d < R * C; ++d) {
    Device &dv = rt.dev(d);
    const Layout &L = plan.layout[size_t(d)];
    unsigned char *B = base[size_t(d)];
    cudaStream_t st = dv.stream;
    CK(cudaSetDevice(dv.id));
    const Readback &rb = readback[size_t(d)];
    if (rb.q > 0) {
      const int *ids_src = rb.folded ? at<int>(B, L.oi) : at<int>(B, L.fid);
      const double *dist_src = rb.folded ? at<double>(B, L.od) : at<double>(B, L.fdist);
      CK(cudaMemcpyAsync(out->label.data() + rb.qb, at<int>(B, L.final_label), sizeof(int) * size_t(rb.q), cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(out->checksum.data() + rb.qb, at<uint64_t>(B, L.final_sum), sizeof(uint64_t) * size_t(rb.q), cudaMemcpyDeviceToHost, st));
      // lists come back dense [q][kmax]; the CSR result is filled after the sync below
      h_ids[size_t(d)].resize(size_t(rb.q) * km);
      h_dist[size_t(d)].resize(size_t(rb.q) * km);
      CK(cudaMemcpyAsync(h_ids[size_t(d)].data(), ids_src, sizeof(int) * size_t(rb.q) * km, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(h_dist[size_t(d)].data(), dist_src, sizeof(double) * size_t(rb.q) * km, cudaMemcpyDeviceToHost, st));
      owned[size_t(d)].qb = rb.qb;
      owned[size_t(d)].q = rb.q;
    }
    CK(cudaMemcpyAsync(&h_status[size_t(d)], at<int>(B, L.status), sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(h_nflag[size_t(d)].data(), at<int>(B, L.nflag), sizeof(int), cudaMemcpyDeviceToHost, st));
  }

  // ---- the one wait of the call: every device's read-back ------------------------------------
  long long total_flagged = 0;
  int worst = 0, bad_status = 0;
  for (int d = 0; d < R * C; ++d) {
    CK(cudaSetDevice(rt.dev(d).id));
    CK(cudaStreamSynchronize(rt.dev(d).stream));
    total_flagged += h_nflag[size_t(d)][0];
    worst = std::max(worst, h_nflag[size_t(d)][0]);
    if (h_status[size_t(d)] != 0) bad_status = h_status[size_t(d)];
  }
  if (bad_status != 0)
    throw std::runtime_error("a device-side wait for a peer GPU timed out (status " + std::to_string(bad_status) + ")");
  if (worst > kRepairCap) {
    // More uncertified queries on one shard than the call repairs by itself (not seen on the
    // generator's distributions): answer the whole job on the direct fp64 kernels instead.
    if (opt.trace)
      for (int d = 0; d < R * C; ++d) {
        std::vector<int> f(size_t(std::min<int64_t>(S.qp, 4096)));
        CK(cudaSetDevice(rt.dev(d).id));
        CK(cudaMemcpy(f.data(), base[size_t(d)] + plan.layout[size_t(d)].flags, f.size() * sizeof(int), cudaMemcpyDeviceToHost));
        int hist[8] = {0};
        for (int v : f) ++hist[v & 7];
        std::fprintf(stderr, "[dmlp] device %d: %d flagged; reason bits (1 cutoff, 2 overflow, 4 model) over the first %zu "
                     "queries: 0:%d 1:%d 2:%d 3:%d 4:%d 5:%d 6:%d 7:%d\n", d, h_nflag[size_t(d)][0], f.size(), hist[0], hist[1],
                     hist[2], hist[3], hist[4], hist[5], hist[6], hist[7]);
      }
    GpuOptions again = opt;
    again.exact_only = true;
    knn_gpu(data, queries, again, out, nullptr);
    if (path) *path += " (re-run exact: " + std::to_string(total_flagged) + " uncertified)";
    return;
  }
  for (int d = 0; d < R * C; ++d) {
    const Owner &o = owned[size_t(d)];
    for (int64_t i = 0; i < o.q; ++i) {
      const int64_t gq = o.qb + i;
      const int kq = int(out->offset[gq + 1] - out->offset[gq]);
      for (int j = 0; j < kq; ++j) {
        out->ids[size_t(out->offset[gq] + j)] = h_ids[size_t(d)][size_t(i) * km + j];
        out->dist[size_t(out->offset[gq] + j)] = h_dist[size_t(d)][size_t(i) * km + j];
      }
    }
  }
  if (opt.trace) {
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::fprintf(stderr, "[dmlp] %s kmax=%d flagged=%lld epoch=%u arena=%.1f MB total=%.3f ms\n",
                 path ? path->c_str() : "gpu", kmax, total_flagged, epoch, plan.bytes / 1048576.0, ms);
  }
}

}  // namespace gpu
}  // namespace dmlp
