"""torch.ops.dmlp_b200.* (src/torch/ops.cpp): registration on any machine, numerics on a GPU."""
import numpy as np
import pytest
import torch

from dmlp_b200.ops import aten, cpu


def test_operators_are_registered_with_schemas():
    ns = aten._ns()
    assert "int kmax" in str(ns.knn_local.default._schema)
    assert "Tensor(a!) dist" in str(ns.exact_.default._schema)
    assert "-> (Tensor, Tensor)" in str(ns.classify.default._schema)
    assert "Tensor?" in str(ns.finalize.default._schema)


def test_cpu_tensors_are_refused_not_silently_computed():
    data = torch.zeros(4, 3, dtype=torch.float64)
    with pytest.raises(NotImplementedError, match="CPU"):
        aten.classify(data, torch.zeros(4, dtype=torch.int32), data[:2].contiguous(), torch.ones(2, dtype=torch.int32), 1)


@pytest.mark.gpu
@pytest.mark.parametrize("n,q,a,kmin,kmax", [(5000, 300, 64, 1, 16), (700, 40, 9, 1, 5), (3000, 260, 128, 20, 40),
                                             (2000, 100, 16, 100, 200)])
def test_classify_matches_oracle(n, q, a, kmin, kmax):
    data, labels, queries, k = cpu.synth(n, q, a, -100.0, 100.0, kmin, kmax, 6, seed=77)
    want_label, want_sum, *_ = cpu.knn(data, labels, queries, k)
    dev = torch.device("cuda:0")
    d, lab = torch.from_numpy(data).to(dev), torch.from_numpy(labels).to(dev)
    qq, kk = torch.from_numpy(queries).to(dev), torch.from_numpy(k).to(dev)
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):                      # the operators follow the current stream
        label, checksum = aten.classify(d, lab, qq, kk, int(k.max()))
    side.synchronize()
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()
    assert (label.cpu().numpy() == want_label).all()


@pytest.mark.gpu
def test_pieces_match_the_ctypes_path():
    from dmlp_b200.ops import cuda as K
    data, labels, queries, k = cpu.synth(6000, 280, 48, -100.0, 100.0, 1, 20, 6, seed=4)
    dev = torch.device("cuda:0")
    d, lab = torch.from_numpy(data).to(dev), torch.from_numpy(labels).to(dev)
    qq, kk = torch.from_numpy(queries).to(dev), torch.from_numpy(k).to(dev)
    kmax = int(k.max())
    got = aten.knn_local(qq, kk, d, lab, kmax, 7)
    ref = K.knn_local(qq, kk, d, lab, kmax, 7)
    torch.cuda.synchronize()
    for g, r in zip(got, ref):
        assert torch.equal(g, r)
    label, checksum, ids, dist = aten.finalize(got[0][None], got[1][None], got[2][None], kk, True)
    rl, rs, ri, rd = K.finalize(ref[0][None], ref[1][None], ref[2][None], kk, True)
    assert torch.equal(label, rl) and torch.equal(checksum, rs) and torch.equal(ids, ri) and torch.equal(dist, rd)
    od, oi, ol = torch.empty_like(got[0]), torch.empty_like(got[1]), torch.empty_like(got[2])
    aten.exact_(qq, d, lab, kk, kmax, 7, torch.empty(0, dtype=torch.int32, device=dev), od, oi, ol)
    torch.cuda.synchronize()
    valid = oi >= 0
    assert torch.equal(oi, got[1]) and torch.equal(od[valid], got[0][valid])
    with pytest.raises(RuntimeError, match="dtype"):
        aten.knn_local(qq.float(), kk, d, lab, kmax)
