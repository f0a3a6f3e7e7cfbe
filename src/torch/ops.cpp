// ATen operator layer: the sm_100a pipeline as torch.ops.dmlp_b200.* on CUDA tensors.
//
// The C ABI (include/dmlp_knn.h) takes raw pointers, a stream and caller-owned scratch.
// These operators are that ABI for PyTorch programs: they validate tensors, run on the
// current CUDA stream of the tensors' device and take outputs and scratch from the
// caching allocator, so they compose with other torch work without synchronising.
//   knn_local(queries, k, data, labels, kmax, id_base) -> (dist, ids, labels, flags, num_flagged)
//   exact_(queries, data, labels, k, kmax, id_base, rows, dist, ids, out_labels)   in place
//   finalize(dist, ids, labels, k, want_lists) -> (label, checksum, ids, dist)
//   classify(data, labels, queries, k, kmax, id_base) -> (label, checksum)
// classify is Engine::KNN for resident tensors (reference: engine.cpp:22-350 end to end):
// filter + certified re-rank, exact fallback for queries the filter could not certify,
// vote and checksum.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include <algorithm>
#include <optional>
#include <tuple>

#include "dmlp_knn.h"

namespace {

using at::Tensor;


void ok(int rc, const char *what) { TORCH_CHECK(rc == 0, what, ": ", dmlp_last_error()); }

void want(const Tensor &t, const char *name, at::ScalarType dtype, int64_t dim) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor");
  TORCH_CHECK(t.scalar_type() == dtype, name, " must have dtype ", dtype, ", got ", t.scalar_type());
  TORCH_CHECK(t.dim() == dim, name, " must have ", dim, " dimensions, got ", t.dim());
  TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
}

void same_device(const Tensor &a, const Tensor &b, const char *what) {
  TORCH_CHECK(a.device() == b.device(), what, " must be on the same device");
}

int sm_count() { return at::cuda::getCurrentDeviceProperties()->multiProcessorCount; }

std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> knn_local(const Tensor &queries, const Tensor &k, const Tensor &data,
                                                             const Tensor &labels, int64_t kmax, int64_t id_base) {
  want(queries, "queries", at::kDouble, 2);
  want(data, "data", at::kDouble, 2);
  want(k, "k", at::kInt, 1);
  want(labels, "labels", at::kInt, 1);
  same_device(queries, data, "queries and data");
  same_device(queries, k, "queries and k");
  same_device(data, labels, "data and labels");
  const int64_t q = queries.size(0), a = queries.size(1), n = data.size(0);
  TORCH_CHECK(data.size(1) == a, "queries have ", a, " attributes, data has ", data.size(1));
  TORCH_CHECK(k.size(0) == q && labels.size(0) == n, "k / labels length does not match the row count");
  TORCH_CHECK(kmax >= 1 && (n == 0 || dmlp_fused_supported(int(kmax), n) >= 1),
              "knn_local supports kmax <= 96, or kmax <= 1024 with n >= 32 kmax; got kmax = ", kmax, ", n = ", n);
  c10::cuda::CUDAGuard guard(queries.device());
  auto stream = at::cuda::getCurrentCUDAStream();
  const auto f64 = queries.options(), i32 = k.options();
  Tensor od = at::empty({q, kmax}, f64), oi = at::empty({q, kmax}, i32), ol = at::empty({q, kmax}, i32);
  Tensor flags = at::zeros({q}, i32), nflag = at::zeros({1}, i32);
  if (q == 0 || n == 0) {
    od.fill_(1e18);
    oi.fill_(-1);
    ol.fill_(-1);
    return {od, oi, ol, flags, nflag};
  }
  const size_t wsb = dmlp_knn_local_workspace_bytes(int(q), int(n), int(a), sm_count(), int(kmax));
  Tensor ws = at::empty({static_cast<int64_t>(wsb)}, queries.options().dtype(at::kByte));
  ok(dmlp_knn_local(queries.data_ptr<double>(), k.data_ptr<int>(), int(q), data.data_ptr<double>(),
                    labels.data_ptr<int>(), int(n), int(a), int(id_base), int(kmax), nullptr, ws.data_ptr(), wsb,
                    od.data_ptr<double>(), oi.data_ptr<int>(), ol.data_ptr<int>(), flags.data_ptr<int>(),
                    nflag.data_ptr<int>(), stream.stream()),
     "knn_local");
  return {od, oi, ol, flags, nflag};
}

// rows: int32 indices of the queries to (re)compute, or an empty tensor for all of them.
void exact_(const Tensor &queries, const Tensor &data, const Tensor &labels, const Tensor &k, int64_t kmax,
            int64_t id_base, const Tensor &rows, Tensor &dist, Tensor &ids, Tensor &out_labels) {
  want(queries, "queries", at::kDouble, 2);
  want(data, "data", at::kDouble, 2);
  want(k, "k", at::kInt, 1);
  want(labels, "labels", at::kInt, 1);
  want(dist, "dist", at::kDouble, 2);
  want(ids, "ids", at::kInt, 2);
  want(out_labels, "out_labels", at::kInt, 2);
  const int64_t q = queries.size(0), a = queries.size(1), n = data.size(0);
  TORCH_CHECK(data.size(1) == a && k.size(0) == q && labels.size(0) == n, "shape mismatch");
  TORCH_CHECK(dist.size(0) == q && dist.size(1) == kmax && ids.sizes() == dist.sizes() &&
                  out_labels.sizes() == dist.sizes(),
              "outputs must be [q, kmax]");
  const bool subset = rows.numel() > 0;
  if (subset) want(rows, "rows", at::kInt, 1);
  const int64_t count = subset ? rows.numel() : q;
  if (count == 0) return;
  c10::cuda::CUDAGuard guard(queries.device());
  auto stream = at::cuda::getCurrentCUDAStream();
  const size_t wsb = dmlp_exact_workspace_bytes(int(n), int(count), int(kmax));
  Tensor ws = at::empty({static_cast<int64_t>(wsb)}, queries.options().dtype(at::kByte));
  ok(dmlp_exact(queries.data_ptr<double>(), data.data_ptr<double>(), labels.data_ptr<int>(), k.data_ptr<int>(),
                subset ? rows.data_ptr<int>() : nullptr, int(count), int(n), int(a), int(id_base), int(kmax),
                dist.data_ptr<double>(), ids.data_ptr<int>(), out_labels.data_ptr<int>(), ws.data_ptr(), wsb,
                stream.stream()),
     "exact");
}

std::tuple<Tensor, Tensor, std::optional<Tensor>, std::optional<Tensor>> finalize(const Tensor &dist, const Tensor &ids,
                                                                                  const Tensor &labels, const Tensor &k,
                                                                                  bool want_lists) {
  want(dist, "dist", at::kDouble, 3);
  want(ids, "ids", at::kInt, 3);
  want(labels, "labels", at::kInt, 3);
  want(k, "k", at::kInt, 1);
  TORCH_CHECK(ids.sizes() == dist.sizes() && labels.sizes() == dist.sizes(), "dist / ids / labels shapes differ");
  const int64_t sources = dist.size(0), q = dist.size(1), kmax = dist.size(2);
  TORCH_CHECK(k.size(0) == q, "k must have one entry per query");
  c10::cuda::CUDAGuard guard(dist.device());
  auto stream = at::cuda::getCurrentCUDAStream();
  Tensor label = at::empty({q}, k.options());
  Tensor checksum = at::empty({q}, k.options().dtype(at::kLong));  // uint64 bit pattern
  Tensor out_id = want_lists ? at::empty({q, kmax}, ids.options()) : Tensor();
  Tensor out_dist = want_lists ? at::empty({q, kmax}, dist.options()) : Tensor();
  if (q > 0) {
    const size_t wsb = dmlp_finalize_workspace_bytes(int(sources), int(q), int(kmax));
    Tensor ws = wsb ? at::empty({static_cast<int64_t>(wsb)}, dist.options().dtype(at::kByte)) : Tensor();
    ok(dmlp_finalize(dist.data_ptr<double>(), ids.data_ptr<int>(), labels.data_ptr<int>(), q * kmax, int(sources),
                     int(q), int(kmax), k.data_ptr<int>(), label.data_ptr<int>(),
                     reinterpret_cast<uint64_t *>(checksum.data_ptr<int64_t>()),
                     want_lists ? out_id.data_ptr<int>() : nullptr, want_lists ? out_dist.data_ptr<double>() : nullptr,
                     wsb ? ws.data_ptr() : nullptr, wsb, stream.stream()),
       "finalize");
  }
  if (!want_lists) return {label, checksum, std::nullopt, std::nullopt};
  return {label, checksum, out_id, out_dist};
}

std::tuple<Tensor, Tensor> classify(const Tensor &data, const Tensor &labels, const Tensor &queries, const Tensor &k,
                                    int64_t kmax, int64_t id_base) {
  TORCH_CHECK(kmax >= 1, "classify needs kmax >= 1");
  const int64_t q = queries.size(0), n = data.size(0), a = queries.dim() == 2 ? queries.size(1) : 0;
  const bool fused = n > 0 && q > 0 && dmlp_fused_supported(int(kmax), n) >= 1 &&
                     double(n) * double(q) * double(std::max<int64_t>(a, 1)) >= double(1 << 22);
  Tensor od, oi, ol;
  if (fused) {
    Tensor flags, nflag;
    std::tie(od, oi, ol, flags, nflag) = knn_local(queries, k, data, labels, kmax, id_base);
    if (nflag.item<int>() != 0) {  // the one host sync of the call
      Tensor rows = at::nonzero(flags).flatten().to(at::kInt);
      exact_(queries, data, labels, k, kmax, id_base, rows, od, oi, ol);
    }
  } else {
    want(queries, "queries", at::kDouble, 2);
    od = at::empty({q, kmax}, queries.options());
    oi = at::empty({q, kmax}, k.options());
    ol = at::empty({q, kmax}, k.options());
    exact_(queries, data, labels, k, kmax, id_base, at::empty({0}, k.options()), od, oi, ol);
  }
  auto res = finalize(od.unsqueeze(0), oi.unsqueeze(0), ol.unsqueeze(0), k, false);
  return {std::get<0>(res), std::get<1>(res)};
}

}  // namespace

TORCH_LIBRARY(dmlp_b200, m) {
  m.def("knn_local(Tensor queries, Tensor k, Tensor data, Tensor labels, int kmax, int id_base=0) -> "
        "(Tensor, Tensor, Tensor, Tensor, Tensor)");
  m.def("exact_(Tensor queries, Tensor data, Tensor labels, Tensor k, int kmax, int id_base, Tensor rows, "
        "Tensor(a!) dist, Tensor(b!) ids, Tensor(c!) out_labels) -> ()");
  m.def("finalize(Tensor dist, Tensor ids, Tensor labels, Tensor k, bool want_lists=False) -> "
        "(Tensor, Tensor, Tensor?, Tensor?)");
  m.def("classify(Tensor data, Tensor labels, Tensor queries, Tensor k, int kmax, int id_base=0) -> (Tensor, Tensor)");
}

TORCH_LIBRARY_IMPL(dmlp_b200, CUDA, m) {
  m.impl("knn_local", &knn_local);
  m.impl("exact_", &exact_);
  m.impl("finalize", &finalize);
  m.impl("classify", &classify);
}
