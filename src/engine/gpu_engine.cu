// GPU backend of Engine::KNNFlat: one host process drives an R x C grid of B200s.
//
// Reference flow being replaced (engine.cpp:40-309): MPI_Dims_create grid,
// Scatterv + Bcast of data shards along rows and query shards along columns,
// local brute force, MPI_Gather of k-lists to row 0, re-select.  Here device
// (r, c) receives data shard r and query shard c straight from (pinned) host
// memory with async copies on its own stream -- no GPU-to-GPU traffic for the
// scatter -- runs the fused local pipeline (dmlp_knn_local) or the direct fp64
// kernel, and ships its exact k-lists to the column owner (0, c) with a peer
// copy; the owner merges, votes and checksums (dmlp_finalize).  Block partitions
// give the remainder to shard 0 like the reference (engine.cpp:62-63, 136-137)
// and non-square grids are handled correctly.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <vector>

#include <cuda_runtime.h>

#include "dmlp/cpu.h"
#include "dmlp/gpu.h"
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

constexpr int kFusedMaxK = 96;  // keep in sync with ops/cuda.py FUSED_MAX_K

struct DeviceBuf {
  void *p = nullptr;
  void alloc(size_t bytes) { CK(cudaMalloc(&p, bytes ? bytes : 256)); }
  void release() { if (p) cudaFree(p); p = nullptr; }
  template <class T> T *as() const { return static_cast<T *>(p); }
};

struct Shard {
  int dev = 0, r = 0, c = 0;
  int64_t n = 0, nb = 0, q = 0, qb = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;
  DeviceBuf x, xl, qa, qk, ws, od, oi, ol, flags, nflag, rows;
  // owner-only
  DeviceBuf gd, gi, gl, label, checksum, fid, fdist;
};

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

void knn_gpu(const DatasetView &data, const QueryView &queries, const GpuOptions &opt, ResultSet *out,
             std::string *path) {
  const int ndev = std::max(1, std::min(opt.ngpu, device_count()));
  int R = opt.rows, C = opt.cols;
  if (R <= 0 || C <= 0 || R * C > ndev) { R = ndev; C = 1; }
  const int a = data.a;
  const int64_t N = data.n, Q = queries.q;

  int kmax = 1;
  for (int64_t i = 0; i < Q; ++i) kmax = std::max(kmax, queries.k[i]);
  const bool fused = !opt.exact_only && kmax <= kFusedMaxK && N > 0 && Q > 0 &&
                     double(N) * double(Q) * double(std::max(a, 1)) >= double(1 << 22);
  const bool device_merge = int64_t(R) * kmax <= 1024;

  out->clear();
  out->offset.assign(size_t(Q) + 1, 0);
  for (int64_t i = 0; i < Q; ++i) out->offset[i + 1] = out->offset[i] + std::max(queries.k[i], 0);
  out->dist.assign(size_t(out->offset[Q]), 0.0);
  out->ids.assign(size_t(out->offset[Q]), -1);
  out->label.assign(size_t(Q), -1);
  out->checksum.assign(size_t(Q), 0);
  if (path) {
    char buf[96];
    std::snprintf(buf, sizeof buf, "gpu:%s R=%d C=%d", fused ? "fused" : "exact", R, C);
    *path = buf;
  }
  if (Q == 0) return;

  BlockPartition dp{N, R}, qp{Q, C};
  std::vector<Shard> shards(size_t(R) * C);
  auto at = [&](int r, int c) -> Shard & { return shards[size_t(r) * C + c]; };
  auto t0 = std::chrono::steady_clock::now();

  // ---- per device: allocate, H2D, local pipeline -------------------------------
  for (int r = 0; r < R; ++r)
    for (int c = 0; c < C; ++c) {
      Shard &s = at(r, c);
      s.dev = r * C + c; s.r = r; s.c = c;
      s.n = dp.count(r); s.nb = dp.begin(r); s.q = qp.count(c); s.qb = qp.begin(c);
      CK(cudaSetDevice(s.dev));
      CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
      CK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
      if (s.q == 0) continue;
      s.x.alloc(sizeof(double) * size_t(s.n) * a);
      s.xl.alloc(sizeof(int) * size_t(s.n));
      s.qa.alloc(sizeof(double) * size_t(s.q) * a);
      s.qk.alloc(sizeof(int) * size_t(s.q));
      s.od.alloc(sizeof(double) * size_t(s.q) * kmax);
      s.oi.alloc(sizeof(int) * size_t(s.q) * kmax);
      s.ol.alloc(sizeof(int) * size_t(s.q) * kmax);
      s.flags.alloc(sizeof(int) * size_t(s.q));
      s.nflag.alloc(sizeof(int));
      CK(cudaMemcpyAsync(s.x.p, data.attrs + s.nb * a, sizeof(double) * size_t(s.n) * a, cudaMemcpyHostToDevice, s.stream));
      CK(cudaMemcpyAsync(s.xl.p, data.labels + s.nb, sizeof(int) * size_t(s.n), cudaMemcpyHostToDevice, s.stream));
      CK(cudaMemcpyAsync(s.qa.p, queries.attrs + s.qb * a, sizeof(double) * size_t(s.q) * a, cudaMemcpyHostToDevice, s.stream));
      CK(cudaMemcpyAsync(s.qk.p, queries.k + s.qb, sizeof(int) * size_t(s.q), cudaMemcpyHostToDevice, s.stream));
      CK(cudaMemsetAsync(s.nflag.p, 0, sizeof(int), s.stream));
      if (fused && s.n > 0) {
        size_t wsb = dmlp_knn_local_workspace_bytes(int(s.q), int(s.n), a, 0, kmax);
        // size with the real SM count of this device
        int sms = 0, ma = 0, mi = 0;
        check_rc(dmlp_device_info(&sms, &ma, &mi), "device_info");
        wsb = dmlp_knn_local_workspace_bytes(int(s.q), int(s.n), a, sms, kmax);
        s.ws.alloc(wsb);
        check_rc(dmlp_knn_local(s.qa.as<double>(), s.qk.as<int>(), int(s.q), s.x.as<double>(), s.xl.as<int>(),
                                int(s.n), a, data.id_base + int(s.nb), kmax, nullptr, s.ws.p, wsb,
                                s.od.as<double>(), s.oi.as<int>(), s.ol.as<int>(), s.flags.as<int>(),
                                s.nflag.as<int>(), s.stream),
                 "knn_local");
      } else {
        size_t wsb = dmlp_exact_workspace_bytes(int(s.n), int(s.q), kmax);
        s.ws.alloc(wsb);
        check_rc(dmlp_exact(s.qa.as<double>(), s.x.as<double>(), s.xl.as<int>(), s.qk.as<int>(), nullptr, int(s.q),
                            int(s.n), a, data.id_base + int(s.nb), kmax, s.od.as<double>(), s.oi.as<int>(),
                            s.ol.as<int>(), s.ws.p, wsb, s.stream),
                 "exact");
      }
    }

  // ---- queries the filter could not certify take the exact kernel --------------
  int64_t total_flagged = 0;
  if (fused)
    for (Shard &s : shards) {
      if (s.q == 0 || s.n == 0) continue;
      CK(cudaSetDevice(s.dev));
      int nflag = 0;
      CK(cudaMemcpyAsync(&nflag, s.nflag.p, sizeof(int), cudaMemcpyDeviceToHost, s.stream));
      CK(cudaStreamSynchronize(s.stream));
      if (nflag == 0) continue;
      total_flagged += nflag;
      std::vector<int> flags(size_t(s.q)), rows;
      CK(cudaMemcpy(flags.data(), s.flags.p, sizeof(int) * size_t(s.q), cudaMemcpyDeviceToHost));
      for (int64_t i = 0; i < s.q; ++i) if (flags[size_t(i)]) rows.push_back(int(i));
      s.rows.alloc(sizeof(int) * rows.size());
      CK(cudaMemcpyAsync(s.rows.p, rows.data(), sizeof(int) * rows.size(), cudaMemcpyHostToDevice, s.stream));
      size_t wsb = dmlp_exact_workspace_bytes(int(s.n), int(rows.size()), kmax);
      DeviceBuf ews; ews.alloc(wsb);
      check_rc(dmlp_exact(s.qa.as<double>(), s.x.as<double>(), s.xl.as<int>(), s.qk.as<int>(), s.rows.as<int>(),
                          int(rows.size()), int(s.n), a, data.id_base + int(s.nb), kmax, s.od.as<double>(),
                          s.oi.as<int>(), s.ol.as<int>(), ews.p, wsb, s.stream),
               "exact(flagged)");
      CK(cudaStreamSynchronize(s.stream));
      ews.release();
    }

  // ---- reduce along the data axis: lists -> column owner, merge, vote ----------
  std::vector<double> h_dist;
  std::vector<int> h_id, h_label_lists;
  for (int c = 0; c < C; ++c) {
    Shard &own = at(0, c);
    if (own.q == 0) continue;
    const size_t per = size_t(own.q) * kmax;
    CK(cudaSetDevice(own.dev));
    own.gd.alloc(sizeof(double) * per * R);
    own.gi.alloc(sizeof(int) * per * R);
    own.gl.alloc(sizeof(int) * per * R);
    for (int r = 0; r < R; ++r) {
      Shard &s = at(r, c);
      CK(cudaSetDevice(s.dev));
      CK(cudaEventRecord(s.done, s.stream));
      CK(cudaSetDevice(own.dev));
      CK(cudaStreamWaitEvent(own.stream, s.done, 0));
      CK(cudaMemcpyPeerAsync(own.gd.as<double>() + per * r, own.dev, s.od.p, s.dev, sizeof(double) * per, own.stream));
      CK(cudaMemcpyPeerAsync(own.gi.as<int>() + per * r, own.dev, s.oi.p, s.dev, sizeof(int) * per, own.stream));
      CK(cudaMemcpyPeerAsync(own.gl.as<int>() + per * r, own.dev, s.ol.p, s.dev, sizeof(int) * per, own.stream));
    }
    if (device_merge) {
      own.label.alloc(sizeof(int) * size_t(own.q));
      own.checksum.alloc(sizeof(uint64_t) * size_t(own.q));
      own.fid.alloc(sizeof(int) * per);
      own.fdist.alloc(sizeof(double) * per);
      check_rc(dmlp_finalize(own.gd.as<double>(), own.gi.as<int>(), own.gl.as<int>(), int64_t(per), R, int(own.q),
                             kmax, own.qk.as<int>(), own.label.as<int>(), own.checksum.as<uint64_t>(),
                             own.fid.as<int>(), own.fdist.as<double>(), nullptr, 0, own.stream),
               "finalize");
      std::vector<int> ids(per);
      std::vector<double> dist(per);
      CK(cudaMemcpyAsync(out->label.data() + own.qb, own.label.p, sizeof(int) * size_t(own.q), cudaMemcpyDeviceToHost, own.stream));
      CK(cudaMemcpyAsync(out->checksum.data() + own.qb, own.checksum.p, sizeof(uint64_t) * size_t(own.q), cudaMemcpyDeviceToHost, own.stream));
      CK(cudaMemcpyAsync(ids.data(), own.fid.p, sizeof(int) * per, cudaMemcpyDeviceToHost, own.stream));
      CK(cudaMemcpyAsync(dist.data(), own.fdist.p, sizeof(double) * per, cudaMemcpyDeviceToHost, own.stream));
      CK(cudaStreamSynchronize(own.stream));
      for (int64_t i = 0; i < own.q; ++i) {
        const int64_t gq = own.qb + i;
        const int kq = int(out->offset[gq + 1] - out->offset[gq]);
        for (int j = 0; j < kq; ++j) {
          out->ids[size_t(out->offset[gq] + j)] = ids[size_t(i) * kmax + j];
          out->dist[size_t(out->offset[gq] + j)] = dist[size_t(i) * kmax + j];
        }
      }
    } else {
      // k beyond the device merge capacity: merge and vote on the host
      h_dist.resize(per * R); h_id.resize(per * R); h_label_lists.resize(per * R);
      CK(cudaMemcpyAsync(h_dist.data(), own.gd.p, sizeof(double) * per * R, cudaMemcpyDeviceToHost, own.stream));
      CK(cudaMemcpyAsync(h_id.data(), own.gi.p, sizeof(int) * per * R, cudaMemcpyDeviceToHost, own.stream));
      CK(cudaMemcpyAsync(h_label_lists.data(), own.gl.p, sizeof(int) * per * R, cudaMemcpyDeviceToHost, own.stream));
      CK(cudaStreamSynchronize(own.stream));
      std::vector<Neighbor> pool;
      for (int64_t i = 0; i < own.q; ++i) {
        const int64_t gq = own.qb + i;
        const int kq = int(out->offset[gq + 1] - out->offset[gq]);
        pool.clear();
        for (int r = 0; r < R; ++r)
          for (int j = 0; j < kq; ++j) {
            size_t o = per * r + size_t(i) * kmax + j;
            pool.push_back(Neighbor{h_dist[o], h_id[o], h_label_lists[o]});
          }
        std::partial_sort(pool.begin(), pool.begin() + kq, pool.end(), closer);
        const int label = vote_label(pool.data(), kq);
        for (int j = 0; j < kq; ++j) {
          out->ids[size_t(out->offset[gq] + j)] = pool[size_t(j)].id;
          out->dist[size_t(out->offset[gq] + j)] = pool[size_t(j)].dist;
        }
        out->label[size_t(gq)] = label;
        out->checksum[size_t(gq)] = result_checksum(label, out->ids.data() + out->offset[gq], kq);
      }
    }
  }

  if (opt.trace) {
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::fprintf(stderr, "[dmlp] %s kmax=%d flagged=%lld total=%.3f ms\n", path ? path->c_str() : "gpu", kmax,
                 (long long)total_flagged, ms);
  }
  for (Shard &s : shards) {
    cudaSetDevice(s.dev);
    for (DeviceBuf *b : {&s.x, &s.xl, &s.qa, &s.qk, &s.ws, &s.od, &s.oi, &s.ol, &s.flags, &s.nflag, &s.rows, &s.gd,
                         &s.gi, &s.gl, &s.label, &s.checksum, &s.fid, &s.fdist})
      b->release();
    if (s.done) cudaEventDestroy(s.done);
    if (s.stream) cudaStreamDestroy(s.stream);
  }
}

}  // namespace gpu
}  // namespace dmlp
