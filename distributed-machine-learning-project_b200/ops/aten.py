"""torch.ops.dmlp_b200.*: the sm_100a pipeline behind PyTorch's dispatcher (src/torch/ops.cpp).

Same kernels as ``ops.cuda`` (which talks to the C ABI through ctypes); here the operators
are registered with ATen, run on the current CUDA stream of their inputs' device and take
outputs and scratch from the caching allocator.  Inputs: fp64 [rows, a] attribute matrices,
int32 labels / k, all contiguous CUDA tensors.

    label, checksum = aten.classify(data, labels, queries, k, kmax)
"""
from .. import _native


def _ns():
    return _native.torch_ops()


def knn_local(queries, k, data, labels, kmax, id_base=0):
    """Certified filter + exact re-rank for one shard: (dist, ids, labels, flags, num_flagged)."""
    return _ns().knn_local(queries, k, data, labels, int(kmax), int(id_base))


def exact_(queries, data, labels, k, kmax, id_base, rows, dist, ids, out_labels):
    """Direct fp64 brute force into dist / ids / out_labels ([q, kmax]); `rows` (int32) selects
    the queries to compute, an empty tensor means all."""
    _ns().exact_(queries, data, labels, k, int(kmax), int(id_base), rows, dist, ids, out_labels)


def finalize(dist, ids, labels, k, want_lists=False):
    """Merge [sources, q, kmax] exact lists, vote, checksum: (label, checksum, ids|None, dist|None)."""
    return _ns().finalize(dist, ids, labels, k, bool(want_lists))


def classify(data, labels, queries, k, kmax, id_base=0):
    """Engine::KNN on resident tensors: (label [q] int32, checksum [q] int64 = uint64 bits)."""
    return _ns().classify(data, labels, queries, k, int(kmax), int(id_base))
