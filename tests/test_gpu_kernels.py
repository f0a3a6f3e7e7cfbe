"""GPU numerics: every kernel against the CPU oracle / a plain fp32-fp64 PyTorch restatement."""
import numpy as np
import pytest
import torch

from dmlp_b200.ops import cpu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from dmlp_b200.ops import cuda
    return cuda


def _dev(x):
    return torch.from_numpy(x).cuda()


def test_synth_matches_host_twin(K):
    data, labels, queries, k = cpu.synth(300, 40, 37, -100.0, 100.0, 2, 9, 5, seed=9)
    d = K.synth_attrs(300, 37, 0, -100.0, 100.0, 9, 0)
    q = K.synth_attrs(40, 37, 0, -100.0, 100.0, 9, 2)
    assert (d.cpu().numpy() == data).all() and (q.cpu().numpy() == queries).all()
    assert (K.synth_labels(300, 0, 5, 9).cpu().numpy() == labels).all()
    assert (K.synth_k(40, 0, 2, 9, 9).cpu().numpy() == k).all()
    # shards are addressable independently
    part = K.synth_attrs(100, 37, 150, -100.0, 100.0, 9, 0)
    assert (part.cpu().numpy() == data[150:250]).all()


def test_pack_against_torch(K):
    data, *_ = cpu.synth(257, 1, 70, -50.0, 150.0, 1, 1, 2, seed=3)
    x = _dev(data)
    st = K.new_stats()
    K.stats_update(x, st)
    h, n2, en = K.pack(x, st, st)
    torch.cuda.synchronize()
    neg_lo, hi = st[:2].cpu().tolist()
    assert -neg_lo == data.min() and hi == data.max()
    c = 0.5 * data.min() + 0.5 * data.max()
    m = max(data.max() - c, c - data.min())
    s = 2.0 ** (14 - int(np.floor(np.log2(m))))
    xp = (x - c) * s
    ref_h = xp.to(torch.float16)
    assert h.shape == (257, 128)
    # the kernel rounds fp64 -> fp16 directly; torch may round through fp32 first, so
    # compare as "at least as close to the exact value, and almost always equal"
    got = h[:, :70].double()
    assert ((got - xp).abs() <= (ref_h.double() - xp).abs()).all() and (h[:, 70:] == 0).all()
    assert (got != ref_h.double()).float().mean().item() < 1e-3
    ref_h = h[:, :70]
    ref_n2 = (ref_h.double() ** 2).sum(1)
    torch.testing.assert_close(n2.double(), ref_n2, rtol=1e-6, atol=0)
    ref_err = ((xp - ref_h.double()) ** 2).sum(1).sqrt()
    assert (en.double() >= ref_err).all() and (en.double() <= ref_err * 1.001 + 1e-6).all()
    maxima = st.view(torch.float32)[4:6].cpu().tolist()
    assert abs(maxima[0] - ref_n2.sqrt().max().item()) <= 1e-3 * maxima[0]


@pytest.mark.parametrize("q,n,a", [(128, 256, 64), (130, 1000, 100), (300, 5000, 256), (64, 300, 8)])
def test_dist_topk_against_torch(K, q, n, a):
    """Fused GEMM + top-k: each chunk list must equal the 32 smallest scores of its
    column range, scores computed in fp32 from the same fp16 operands."""
    data, _, queries, _ = cpu.synth(n, q, a, -100.0, 100.0, 1, 1, 2, seed=5)
    x, qq = _dev(data), _dev(queries)
    st = K.new_stats()
    K.stats_update(x, st)
    K.stats_update(qq, st)
    xh, xn, _ = K.pack(x, st, st)
    qh, _, _ = K.pack(qq, st)
    cand = K.dist_topk(qh, xh, xn)
    torch.cuda.synchronize()
    plan = K.topk_plan(q, n, xh.shape[1])
    chunks, nt = plan["chunks"], plan["n_tiles"]
    assert cand.shape == (q, chunks, 32)
    scores = xn[None, :].double() - 2.0 * (qh.double() @ xh.double().T)   # [q, n], fp64 reference
    keys = cand.cpu().numpy().view(np.uint64)
    ids = (keys & np.uint64(0xFFFFFFFF)).astype(np.int64)
    valid = keys != np.uint64(0xFFFFFFFFFFFFFFFF)
    sc = scores.cpu().numpy()
    mag = np.abs(sc).max()
    for c in range(chunks):
        lo, hi = (c * nt // chunks) * 256, min(((c + 1) * nt // chunks) * 256, n)
        want_cnt = min(32, hi - lo)
        assert (valid[:, c, :].sum(1) == want_cnt).all()
        for r in range(0, q, max(1, q // 16)):
            got = ids[r, c, :want_cnt]
            assert ((got >= lo) & (got < hi)).all() and len(set(got.tolist())) == want_cnt
            kth = np.partition(sc[r, lo:hi], want_cnt - 1)[want_cnt - 1]
            # everything kept is within fp32 accumulation error of the true 32 best
            assert (sc[r, got] <= kth + 1e-5 * mag).all()
            assert (np.diff(sc[r, got]) >= -1e-5 * mag).all()  # emitted ascending


@pytest.mark.parametrize("q,n,a,chunks,gmc,sync", [(600, 5000, 64, 3, 2, 0), (300, 5000, 256, 2, 1, 0),
                                                   (1000, 9000, 128, 5, 2, 0), (257, 777, 128, 0, 0, 0),
                                                   (1000, 5000, 1024, 3, 2, 1), (1000, 5000, 1024, 3, 2, 3)])
def test_dist_topk_wide_tile_equals_narrow(K, q, n, a, chunks, gmc, sync, monkeypatch):
    """The wide (256 x 512 per pair) kernel must produce the very lists of the narrow one.
    The forced decompositions give chunks of 6/7/7, 10/10 and 7/7/7/7/8 tiles (odd counts end
    with a step that fills one accumulator only) and several items per cluster (the
    accumulator bookkeeping carries over from item to item)."""
    if chunks:
        monkeypatch.setenv("DMLP_TOPK_CHUNKS", str(chunks))
        monkeypatch.setenv("DMLP_TOPK_GMC", str(gmc))
    if sync:  # re-alignment of the pairs that sweep a chunk, every `sync` tiles (both kernels)
        monkeypatch.setenv("DMLP_TOPK_SYNC", str(sync))
    monkeypatch.setenv("DMLP_TOPK_HALVES", "1")   # one list per (query, chunk) in both, like the wide tile
    data, _, queries, _ = cpu.synth(n, q, a, -100.0, 100.0, 1, 1, 2, seed=11)
    x, qq = _dev(data), _dev(queries)
    st = K.new_stats()
    K.stats_update(x, st)
    K.stats_update(qq, st)
    xh, xn, _ = K.pack(x, st, st)
    qh, _, _ = K.pack(qq, st)
    got = {}
    # narrow; wide in lock step (the default); wide skewed (opt-in, two gaps)
    for name, wide, gap in (("narrow", "0", None), ("lockstep", "1", None), ("skew", "1", "12"), ("skew4", "1", "4")):
        monkeypatch.setenv("DMLP_TOPK_WIDE", wide)
        if gap is None:
            monkeypatch.delenv("DMLP_TOPK_GAP", raising=False)
        else:
            monkeypatch.setenv("DMLP_TOPK_GAP", gap)
        got[name] = K.dist_topk(qh, xh, xn).clone()
        torch.cuda.synchronize()
    assert got["narrow"].shape == got["lockstep"].shape and torch.equal(got["narrow"], got["lockstep"])
    # the skewed schedule walks the k-blocks cyclically from a tile-dependent start, so fp32
    # accumulation order -- and the last bits of a score -- differ: compare what matters, the
    # certified exact lists the re-rank derives from the candidates
    kk = torch.full((q,), 9, dtype=torch.int32, device="cuda")
    labels = torch.zeros(n, dtype=torch.int32, device="cuda")
    qst = K.new_stats()
    _, qn, qe = K.pack(qq, st, qst)
    ref = None
    for name in ("narrow", "skew", "skew4"):
        od, oi, ol, flags, nflag = K.rerank(got[name], qq, x, labels, kk, qn, qe, st, st, 0, 9)
        torch.cuda.synchronize()
        assert int(nflag.item()) == 0, name
        if ref is None:
            ref = (od.clone(), oi.clone())
        else:
            assert torch.equal(ref[0], od) and torch.equal(ref[1], oi), name


@pytest.mark.parametrize("q,n,a,chunks,gmc", [(600, 5000, 64, 3, 2), (1000, 9000, 128, 5, 2), (257, 777, 128, 0, 0),
                                              (1000, 5000, 1024, 3, 2), (5000, 20000, 64, 0, 0),
                                              (40000, 3000, 64, 4, 3)])
def test_dist_topk_dynamic_schedule_equals_static(K, q, n, a, chunks, gmc, monkeypatch):
    """DMLP_TOPK_DYN=1: clusters take their items from an atomic counter (the leader's producer
    thread grabs, every other role reads the cluster's slot).  Which cluster computes a (query block,
    chunk) list does not change the list: the candidates must be bit-identical to the static
    round-robin, for the narrow and the wide tile, with few and with many items per cluster."""
    if chunks:
        monkeypatch.setenv("DMLP_TOPK_CHUNKS", str(chunks))
        monkeypatch.setenv("DMLP_TOPK_GMC", str(gmc))
    data, _, queries, _ = cpu.synth(n, q, a, -100.0, 100.0, 1, 1, 2, seed=12)
    x, qq = _dev(data), _dev(queries)
    st = K.new_stats()
    K.stats_update(x, st)
    K.stats_update(qq, st)
    xh, xn, _ = K.pack(x, st, st)
    qh, _, _ = K.pack(qq, st)
    for wide in ("0", "1"):
        monkeypatch.setenv("DMLP_TOPK_WIDE", wide)
        monkeypatch.setenv("DMLP_TOPK_DYN", "0")
        ref = K.dist_topk(qh, xh, xn).clone()
        monkeypatch.setenv("DMLP_TOPK_DYN", "1")
        for _ in range(2):      # twice: the scheduler's scratch is handed out and zeroed per launch
            got = K.dist_topk(qh, xh, xn).clone()
            torch.cuda.synchronize()
            assert torch.equal(ref, got), wide


@pytest.mark.parametrize("q,n,a,chunks,gmc", [(600, 5000, 64, 3, 2), (1000, 9000, 128, 5, 2), (257, 777, 100, 0, 0),
                                              (5000, 30000, 64, 0, 0), (3000, 2000, 32, 2, 3)])
def test_dist_topk_two_epilogue_sets_for_short_rows(K, q, n, a, chunks, gmc, monkeypatch):
    """DMLP_TOPK_HALVES=2 (the default for rows of one k-block): two sets of epilogue warps, each scanning half of a tile's columns
    into lists of its own.  Every (query, chunk) then yields two lists whose union holds the single
    list of the one-set kernel; the re-rank must derive identical exact lists from either."""
    if chunks:
        monkeypatch.setenv("DMLP_TOPK_CHUNKS", str(chunks))
        monkeypatch.setenv("DMLP_TOPK_GMC", str(gmc))
    data, _, queries, _ = cpu.synth(n, q, a, -100.0, 100.0, 1, 1, 2, seed=14)
    x, qq = _dev(data), _dev(queries)
    st, qst = K.new_stats(), K.new_stats()
    K.stats_update(x, st)
    K.stats_update(qq, st)
    xh, xn, _ = K.pack(x, st, st)
    qh, qn, qe = K.pack(qq, st, qst)
    monkeypatch.setenv("DMLP_TOPK_HALVES", "1")
    one = K.dist_topk(qh, xh, xn).clone()
    monkeypatch.setenv("DMLP_TOPK_HALVES", "2")
    two = K.dist_topk(qh, xh, xn).clone()
    torch.cuda.synchronize()
    assert two.shape[1] == 2 * one.shape[1]
    # a key kept by the one-set kernel is among the 32 best of its chunk, hence of its half
    a1 = one.cpu().numpy().view(np.uint64)
    a2 = two.cpu().numpy().view(np.uint64)
    for row in (0, q // 2, q - 1):
        for c in range(one.shape[1]):
            kept = set(a1[row, c].tolist())
            both = set(a2[row, 2 * c].tolist()) | set(a2[row, 2 * c + 1].tolist())
            assert kept <= both, (row, c)
    kk = torch.full((q,), 9, dtype=torch.int32, device="cuda")
    labels = torch.zeros(n, dtype=torch.int32, device="cuda")
    ref = None
    for cand in (one, two):
        od, oi, ol, flags, nflag = K.rerank(cand, qq, x, labels, kk, qn, qe, st, st, 0, 9)
        torch.cuda.synchronize()
        assert int(nflag.item()) == 0
        if ref is None:
            ref = (od.clone(), oi.clone())
        else:
            assert torch.equal(ref[0], od) and torch.equal(ref[1], oi)


def test_knn_local_long_rows_take_the_wide_tile(K):
    # 4096 values per row: the planner picks the wide tile by itself
    n, q, a, kmax = 3000, 520, 4096, 12
    data, labels, queries, k = cpu.synth(n, q, a, -100.0, 100.0, 3, kmax, 5, seed=21)
    want_label, want_sum, *_ = cpu.knn(data, labels, queries, k)
    od, oi, ol, flags, nflag = K.knn_local(_dev(queries), _dev(k), _dev(data), _dev(labels), kmax)
    label, checksum = K.finalize(od[None], oi[None], ol[None], _dev(k))[:2]
    torch.cuda.synchronize()
    assert int(nflag.item()) == 0
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()
    assert (label.cpu().numpy() == want_label).all()


@pytest.mark.parametrize("n,q,a,kmin,kmax,lo,hi", [
    (5000, 300, 64, 1, 16, -100.0, 100.0), (2049, 257, 33, 3, 24, 0.0, 10.0), (900, 129, 7, 1, 5, -1.0, 1.0),
    (40000, 256, 128, 16, 16, -100.0, 100.0), (300, 130, 256, 1, 8, 0.0, 1000.0),
    (30000, 300, 64, 30, 60, -100.0, 100.0), (8000, 150, 32, 96, 96, 0.0, 10.0)])
def test_knn_local_matches_oracle(K, n, q, a, kmin, kmax, lo, hi):
    data, labels, queries, k = cpu.synth(n, q, a, lo, hi, kmin, kmax, 7, seed=13)
    want_label, want_sum, want_ids, want_dist, off = cpu.knn(data, labels, queries, k)
    od, oi, ol, flags, nflag = K.knn_local(_dev(queries), _dev(k), _dev(data), _dev(labels), kmax)
    label, checksum, ids, dist = K.finalize(od[None], oi[None], ol[None], _dev(k), True)
    torch.cuda.synchronize()
    assert int(nflag.item()) == 0, f"filter failed to certify {int(nflag.item())} queries"
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()
    assert (label.cpu().numpy() == want_label).all()
    got_ids, got_dist = ids.cpu().numpy(), dist.cpu().numpy()
    for i in range(q):
        assert (got_ids[i, :k[i]] == want_ids[off[i]:off[i + 1]]).all()
        assert (got_dist[i, :k[i]] == want_dist[off[i]:off[i + 1]]).all()   # bit-identical fp64


@pytest.mark.parametrize("n,q,a,kmin,kmax", [(12000, 130, 24, 97, 128), (40000, 300, 32, 200, 300),
                                             (40000, 140, 16, 900, 1024), (70000, 260, 64, 1, 600)])
def test_fused_path_large_k_matches_oracle(K, n, q, a, kmin, kmax):
    """96 < k <= 1024 on the filter + certified re-rank path: one 32-key list per 8 of k, so k > 256
    runs the filter as several launches over row ranges of the data (plan `parts`); the re-rank
    keeps about 2 k survivors per query.  Bit-identical lists, no query left to the direct path."""
    import time
    assert K.fused_supported(kmax, n) and not K.fused_supported(kmax, n, gather=True)
    assert not K.fused_supported(1025, 10 ** 7) and not K.fused_supported(1024, 32512) and K.fused_supported(1024, 32513)
    assert not K.fused_supported(300, 16128) and K.fused_supported(300, 16129) and not K.fused_supported(128, 3840)
    data, labels, queries, k = cpu.synth(n, q, a, -100.0, 100.0, kmin, kmax, 7, seed=17)
    want_label, want_sum, want_ids, want_dist, off = cpu.knn(data, labels, queries, k)
    plan = K.topk_plan(q, n, K.pad64(a), kmax=kmax)
    assert plan["chunks"] * 8 >= kmax
    qd, kd, xd, ld = _dev(queries), _dev(k), _dev(data), _dev(labels)
    K.knn_local(qd, kd, xd, ld, kmax)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    od, oi, ol, flags, nflag = K.knn_local(qd, kd, xd, ld, kmax)
    label, checksum, ids, dist = K.finalize(od[None], oi[None], ol[None], kd, True)
    torch.cuda.synchronize()
    print(f"fused n={n} q={q} a={a} k<={kmax}: {(time.perf_counter() - t0) * 1e3:.2f} ms, lists {plan['chunks']}")
    assert int(nflag.item()) == 0, f"filter failed to certify {int(nflag.item())} queries"
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()
    assert (label.cpu().numpy() == want_label).all()
    got_ids, got_dist = ids.cpu().numpy(), dist.cpu().numpy()
    for i in range(q):
        assert (got_ids[i, :k[i]] == want_ids[off[i]:off[i + 1]]).all()
        assert (got_dist[i, :k[i]] == want_dist[off[i]:off[i + 1]]).all()   # bit-identical fp64


def test_engine_binary_large_k_single_gpu(repo_root):
    """`./engine` on one GPU with k up to 400: the C++ runtime takes the fused path too, output equal to its CPU backend."""
    import json
    import os
    import subprocess
    exe = os.path.join(repo_root, "engine")
    if not os.path.exists(exe):
        pytest.skip("engine not built")
    args = [exe, "--json", "--synth", "40000", "300", "24", "-100", "100", "200", "400", "5", "9"]

    def run(backend):
        res = subprocess.run(args, capture_output=True, text=True, timeout=240,
                             env=dict(os.environ, DMLP_BACKEND=backend, DMLP_NGPU="1"))
        assert res.returncode == 0, res.stderr[-2000:]
        info = json.loads([ln for ln in res.stderr.splitlines() if ln.startswith("{")][-1])
        return sorted(res.stdout.splitlines()), info

    want, _ = run("cpu")
    got, info = run("gpu")
    assert "fused" in info["backend"], info
    assert got == want


def test_classifier_takes_fused_path_for_large_k(K):
    from dmlp_b200.models import KNNClassifier, choose_path
    assert choose_path(40000, 500, 32, 1000) == "fused" and choose_path(20000, 500, 32, 1000) == "exact"
    assert choose_path(40000, 500, 32, 1025) == "exact"
    data, labels, queries, k = cpu.synth(40000, 150, 20, -100.0, 100.0, 300, 1000, 4, seed=19)
    _, want_sum, *_ = cpu.knn(data, labels, queries, k)
    res = KNNClassifier().fit(data, labels).classify(queries, k)
    assert res.path.startswith("fused"), res.path
    assert (res.checksums_u64() == want_sum).all()


@pytest.mark.parametrize("n,q,a,kmin,kmax", [(1000, 50, 5, 1, 70), (37, 20, 3, 37, 37), (5000, 33, 16, 100, 300)])
def test_exact_path_matches_oracle(K, n, q, a, kmin, kmax):
    data, labels, queries, k = cpu.synth(n, q, a, -5.0, 5.0, kmin, kmax, 3, seed=21)
    want_label, want_sum, *_ = cpu.knn(data, labels, queries, k)
    km = int(k.max())
    od, oi, ol = K.exact(_dev(queries), _dev(data), _dev(labels), _dev(k), km)
    label, checksum, _, _ = K.finalize(od[None], oi[None], ol[None], _dev(k))
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()
    assert (label.cpu().numpy() == want_label).all()


@pytest.mark.parametrize("n,q,a,k_lo,k_hi", [(20000, 64, 32, 250, 300), (9000, 40, 16, 900, 1000),
                                             (3000, 5, 8, 2500, 3000), (70000, 3, 24, 4, 9)])
def test_exact_streaming_large_k_and_splits(K, n, q, a, k_lo, k_hi):
    """k = 300 / 1000 / 3000 and a handful of queries over many rows (data splits + rank merge):
    the streaming kernel keeps bounded per-query lists, no distance matrix, no library sort."""
    import time
    data, labels, queries, k = cpu.synth(n, q, a, -100.0, 100.0, k_lo, k_hi, 6, seed=29)
    want_label, want_sum, want_ids, want_dist, off = cpu.knn(data, labels, queries, k)
    km = int(k.max())
    xd, ld, qd, kd = _dev(data), _dev(labels), _dev(queries), _dev(k)
    K.exact(qd, xd, ld, kd, km)                      # warm-up (module load)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    od, oi, ol = K.exact(qd, xd, ld, kd, km)
    label, checksum, ids, dist = K.finalize(od[None], oi[None], ol[None], kd, True)
    torch.cuda.synchronize()
    print(f"exact+finalize n={n} q={q} a={a} k<={km}: {(time.perf_counter() - t0) * 1e3:.2f} ms")
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()
    assert (label.cpu().numpy() == want_label).all()
    got_ids, got_dist = ids.cpu().numpy(), dist.cpu().numpy()
    for i in range(q):
        assert (got_ids[i, :k[i]] == want_ids[off[i]:off[i + 1]]).all()
        assert (got_dist[i, :k[i]] == want_dist[off[i]:off[i + 1]]).all()


def test_exact_subset_rows_and_small_workspace(K):
    """query_rows subset written at the queries' own rows; a workspace sized for one query
    forces the launcher into batches."""
    data, labels, queries, k = cpu.synth(6000, 90, 12, -10.0, 10.0, 1, 40, 4, seed=33)
    _, want_sum, want_ids, _, off = cpu.knn(data, labels, queries, k)
    km = int(k.max())
    xd, ld, qd, kd = _dev(data), _dev(labels), _dev(queries), _dev(k)
    rows = torch.tensor([3, 77, 41, 0, 89], dtype=torch.int32, device="cuda")
    od = torch.full((90, km), -5.0, dtype=torch.float64, device="cuda")
    oi = torch.full((90, km), -5, dtype=torch.int32, device="cuda")
    ol = torch.full((90, km), -5, dtype=torch.int32, device="cuda")
    K.exact(qd, xd, ld, kd, km, 0, rows, out=(od, oi, ol))
    torch.cuda.synchronize()
    touched = set(rows.cpu().tolist())
    for i in range(90):
        if i in touched:
            assert (oi[i, :k[i]].cpu().numpy() == want_ids[off[i]:off[i + 1]]).all()
            assert (oi[i, k[i]:] == -1).all()
        else:
            assert (oi[i] == -5).all()
    small = torch.empty(K.exact_workspace_bytes(6000, 1, km), dtype=torch.uint8, device="cuda")
    od2, oi2, ol2 = K.exact(qd, xd, ld, kd, km, workspace=small)
    _, checksum, _, _ = K.finalize(od2[None], oi2[None], ol2[None], kd)
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()


@pytest.mark.parametrize("shards,kmin,kmax", [(8, 200, 300), (3, 2100, 2300)])
def test_finalize_merges_long_lists(K, shards, kmin, kmax):
    """sources * kmax far beyond the old 1024-candidate staging buffer (2400, 6900): the merge
    ranks by binary search; lists longer than 2048 use the scratch workspace."""
    n = 9000
    data, labels, queries, k = cpu.synth(n, 24, 10, -100.0, 100.0, kmin, kmax, 5, seed=41)
    _, want_sum, *_ = cpu.knn(data, labels, queries, k)
    km = int(k.max())
    qd, kd = _dev(queries), _dev(k)
    bounds = [n * s // shards for s in range(shards + 1)]
    parts = [K.exact(qd, _dev(data[lo:hi]), _dev(labels[lo:hi]), kd, km, lo) for lo, hi in zip(bounds, bounds[1:])]
    dist, ids, labs = (torch.stack([p[j] for p in parts]) for j in range(3))
    _, checksum, _, _ = K.finalize(dist, ids, labs, kd)
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()


@pytest.mark.parametrize("n,q,a,kmax", [(5000, 300, 64, 16), (2049, 257, 1030, 9)])
def test_decimal6_rows_give_identical_pack_rerank_and_exact(K, n, q, a, kmax):
    """Resident decimal6 rows (int32 i, value i / 1e6): pack, the exact re-rank and the direct kernel
    must produce, bit for bit, what they produce from the fp64 rows."""
    data, labels, queries, k = cpu.synth(n, q, a, -100.0, 100.0, 1, kmax, 5, seed=19)
    x, qq, lab, kk = _dev(data), _dev(queries), _dev(labels), _dev(k)
    xi, ok = K.encode_decimal6(x)
    assert ok and torch.equal(K.expand_decimal6(xi, torch.empty_like(x)), x)
    st = K.new_stats()
    K.stats_update(x, st)
    K.stats_update(qq, st)
    st2 = st.clone()
    h0, n0, e0 = K.pack(x, st, st)
    h1, n1, e1 = K.pack_i32(xi, st2, st2)
    torch.cuda.synchronize()
    assert torch.equal(h0, h1) and torch.equal(n0, n1) and torch.equal(e0, e1)
    assert torch.equal(st.view(torch.int64)[:3], st2.view(torch.int64)[:3])   # range and row maxima (bit patterns)
    qst = K.new_stats()
    qh, qn, qe = K.pack(qq, st, qst)
    cand = K.dist_topk(qh, h0, n0, kmax=kmax)
    ref = K.rerank(cand, qq, x, lab, kk, qn, qe, st, st, 0, kmax)
    dev = x.device
    out = (torch.empty_like(ref[0]), torch.empty_like(ref[1]), torch.empty_like(ref[2]))
    flags, nflag = torch.empty(q, dtype=torch.int32, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
    K.rerank_ex(cand, qq, None, lab, kk, qn, qe, st, st, 0, kmax, flags, nflag, out=out, x_i32=xi)
    torch.cuda.synchronize()
    assert int(nflag.item()) == 0 and int(ref[4].item()) == 0
    for r, o in zip(ref[:3], out):
        assert torch.equal(r, o)
    e_ref = K.exact(qq, x, lab, kk, kmax)
    e_i32 = K.exact(qq, None, lab, kk, kmax, x_i32=xi)
    for r, o in zip(e_ref, e_i32):
        assert torch.equal(r, o)


def test_duplicates_tie_break_on_gpu(K):
    rng = np.random.default_rng(0)
    base = rng.uniform(-1, 1, (50, 16))
    data = np.concatenate([base, base, base])          # every point three times
    labels = rng.integers(0, 4, 150).astype(np.int32)
    queries = rng.uniform(-1, 1, (140, 16))
    k = np.full(140, 7, np.int32)
    _, want_sum, *_ = cpu.knn(data, labels, queries, k)
    for fn in ("fused", "exact"):
        if fn == "fused":
            od, oi, ol, flags, nflag = K.knn_local(_dev(queries), _dev(k), _dev(data), _dev(labels), 7)
            rows = torch.nonzero(flags).flatten().to(torch.int32)
            if rows.numel():
                K.exact(_dev(queries), _dev(data), _dev(labels), _dev(k), 7, 0, rows, out=(od, oi, ol))
        else:
            od, oi, ol = K.exact(_dev(queries), _dev(data), _dev(labels), _dev(k), 7)
        _, checksum, _, _ = K.finalize(od[None], oi[None], ol[None], _dev(k))
        assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all(), fn


def test_finalize_merges_shards(K):
    data, labels, queries, k = cpu.synth(3000, 200, 24, -100.0, 100.0, 2, 12, 5, seed=17)
    _, want_sum, *_ = cpu.knn(data, labels, queries, k)
    parts = []
    for lo, hi in [(0, 1100), (1100, 1900), (1900, 3000)]:
        od, oi, ol, flags, nflag = K.knn_local(_dev(queries), _dev(k), _dev(data[lo:hi]), _dev(labels[lo:hi]), 12, lo)
        assert int(nflag.item()) == 0
        parts.append((od, oi, ol))
    dist = torch.stack([p[0] for p in parts])
    ids = torch.stack([p[1] for p in parts])
    labs = torch.stack([p[2] for p in parts])
    _, checksum, _, _ = K.finalize(dist, ids, labs, _dev(k))
    assert (checksum.cpu().numpy().view(np.uint64) == want_sum).all()


def test_model_api_all_paths(K):
    from dmlp_b200.models import knn
    data, labels, queries, k = cpu.synth(6000, 300, 48, -100.0, 100.0, 1, 20, 6, seed=4)
    _, want_sum, *_ = cpu.knn(data, labels, queries, k)
    for mode in ("auto", "fused", "exact"):
        res = knn(data, labels, queries, k, mode=mode)
        assert (res.checksums_u64() == want_sum).all(), mode
    big_k = np.full(300, 64, np.int32)
    _, want_sum, *_ = cpu.knn(data, labels, queries, big_k)
    assert (knn(data, labels, queries, big_k).checksums_u64() == want_sum).all()


def test_model_update_records(K):
    """Update{id, new_attrs} (reference: common.h:35-38, parse_update common.cpp:46-55) applied
    to the resident dataset: point 17 moves onto query 5 and becomes its nearest neighbour."""
    from dmlp_b200.models import KNNClassifier
    data, labels, queries, k = cpu.synth(3000, 200, 32, -100.0, 100.0, 1, 12, 5, seed=31)
    model = KNNClassifier(want_lists=True).fit(data, labels)
    ident, vals = cpu.parse_update("17 " + " ".join(f"{v:.6f}" for v in queries[5]))
    assert ident == 17 and (vals == queries[5]).all()
    model.update(ident, vals, new_labels=4)
    data2, labels2 = data.copy(), labels.copy()
    data2[17], labels2[17] = vals, 4
    _, want_sum, *_ = cpu.knn(data2, labels2, queries, k)
    res = model.classify(queries, k)
    assert (res.checksums_u64() == want_sum).all()
    assert int(res.ids[5, 0]) == 17 and float(res.dist[5, 0]) == 0.0
    with pytest.raises(IndexError):
        model.update(3000, vals)
    with pytest.raises(ValueError):
        model.update(1, vals[:-1])


def test_decimal6_wire_round_trip(K):
    data, *_ = cpu.synth(500, 1, 33, -100.0, 100.0, 1, 1, 2, seed=8)
    x = _dev(data)
    enc, ok = K.encode_decimal6(x)
    assert ok
    out = K.expand_decimal6(enc, torch.empty_like(x))
    assert torch.equal(out, x)                                  # bit-exact doubles back
    assert (enc.cpu().numpy() == np.rint(data * 1e6).astype(np.int64)).all()
    bad = x.clone()
    bad[3, 3] = 0.1234567                                       # seven decimals: not representable
    assert not K.encode_decimal6(bad)[1]
    big = x.clone()
    big[0, 0] = 5000.0                                          # 5e9 does not fit 32 bits
    assert not K.encode_decimal6(big)[1]


@pytest.mark.parametrize("a,bits", [(64, 28), (33, 28), (5, 9), (100, 31), (7, 2), (64, 32)])
def test_bit_packed_wire_round_trip(K, a, bits):
    """pack_bits / unpack_bits against a host restatement of the wire format: value i travels as
    the `bits`-bit number i + 2^(bits-1), rows are little-endian bit streams padded to 32-bit words."""
    rng = np.random.default_rng(a * 100 + bits)
    half = 1 << (bits - 1)
    vals = rng.integers(-half, half, size=(301, a), dtype=np.int64)
    vals[0, :] = -half
    vals[1, :] = half - 1
    src = torch.from_numpy(vals.astype(np.int32)).cuda()
    words = K.pack_bits(src, bits)
    assert words.shape == (301, K.wire_words(a, bits))
    # host restatement
    u = (vals + half).astype(object)
    expect = np.zeros((301, K.wire_words(a, bits)), dtype=np.uint32)
    for r in (0, 1, 2, 150, 300):
        stream = 0
        for c in range(a):
            stream |= int(u[r, c]) << (c * bits)
        for w in range(expect.shape[1]):
            expect[r, w] = (stream >> (32 * w)) & 0xFFFFFFFF
        assert (words[r].cpu().numpy().view(np.uint32) == expect[r]).all(), r
    back = K.unpack_bits(words, a, bits, torch.empty_like(src))
    assert torch.equal(back, src)
    if bits <= 31:
        dbl = K.unpack_bits(words, a, bits, torch.empty((301, a), dtype=torch.float64, device="cuda"))
        assert torch.equal(dbl, K.expand_decimal6(src, torch.empty_like(dbl)))


@pytest.mark.parametrize("n_total,parts", [(3000, 1), (9000, 2), (20000, 4)])
def test_streamed_end_to_end_single_gpu(K, n_total, parts, monkeypatch):
    """run_from_host: copies overlapped with compute; with a large enough shard the data is
    cut into parts and every (query chunk, data part) pair is filtered as it arrives."""
    from dmlp_b200.parallel import DistributedKNN
    eng = DistributedKNN(device=torch.device("cuda", 0))
    eng.synth_problem(n_total=n_total, q_total=40000, a=64, lo=-100.0, hi=100.0, min_k=1, max_k=12, num_labels=5,
                      seed=3)
    _, ref = eng.run_resident()
    ref = ref.cpu().numpy().copy()
    for wire in ("auto", "int32", "fp64"):       # bit-packed decimal6, plain int32 decimal6, doubles
        monkeypatch.setenv("DMLP_WIRE_BITS", "0" if wire == "int32" else "auto")
        eng.stage_host_inputs(wire="auto" if wire == "int32" else wire)
        assert eng.streamed and eng.wire == ("fp64" if wire == "fp64" else "decimal6") and eng.x_parts == parts
        assert eng.wire_bits == (28 if wire == "auto" else 0)
        for step in range(4):      # steps 0 and 1 also start the next step's copies (second wire set)
            _, chk = eng.run_from_host(prefetch_next=step < 2)
            assert (chk.numpy() == ref).all(), wire
    assert eng.verify_against_oracle(200) == 200
