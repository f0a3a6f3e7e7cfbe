"""Work decomposition of the filter GEMM, checked on the host (no GPU needed).

The planner (make_topk_plan) and the kernel's item enumeration (get_item) are a
contract: every (128-query row block, 256-point data tile) must be visited exactly
once, every candidate list (row block, chunk) must be produced exactly once, and the
clusters that re-align on a (wave group, chunk) counter must all be resident at the
same time.  dmlp_debug_topk_items replays the enumeration on the host.
"""
import ctypes as C

import numpy as np
import pytest

from dmlp_b200 import _native

SMS = 148


def _lib():
    try:
        lib = _native.knn_lib()
    except OSError as e:  # CPU-only build of the repo (make NO_CUDA=1)
        pytest.skip(f"libdmlp_knn.so not built: {e}")
    lib.dmlp_topk_plan.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int * 8)]
    lib.dmlp_topk_plan_ex.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int * 16)]
    lib.dmlp_debug_topk_items.argtypes = [C.c_int] * 5 + [C.c_void_p, C.c_int]
    return lib


def plan_and_items(q, n, a_pad, kmax, sms=SMS):
    lib = _lib()
    arr = (C.c_int * 16)()
    assert lib.dmlp_topk_plan_ex(q, n, a_pad, sms, kmax, C.byref(arr)) == 0
    keys = ["m_blocks", "n_tiles", "group_m", "lists", "k_blocks", "grid", "kp", "cluster", "halves", "parts",
            "chunks", "wide"]
    plan = dict(zip(keys, list(arr)))
    assert plan["lists"] == plan["chunks"] * plan["halves"] * plan["parts"]
    assert plan["halves"] == (2 if (plan["k_blocks"] == 1 and plan["m_blocks"] >= 2) else 1)
    cap = 1 << 16
    buf = np.zeros((cap, 8), dtype=np.int32)
    cnt = lib.dmlp_debug_topk_items(q, n, a_pad, sms, kmax, buf.ctypes.data, cap)
    assert 0 < cnt <= cap
    return plan, buf[:cnt]


SHAPES = [
    (65536, 65536, 8192, 16),     # headline: wide pair kernel, 37 x 2
    (16384, 262144, 1024, 16),    # mid: narrow pair kernel
    (10000, 1000000, 64, 16),     # short rows, odd tile counts per chunk
    (9000, 777, 128, 4),          # fewer tiles than clusters
    (300, 5000, 64, 8),           # three row blocks
    (128, 100000, 64, 16),        # a single row block: one-CTA kernel
    (65536, 65536, 8192, 96),     # large k: many chunks
    (1, 1, 64, 1),
    (257, 257, 4096, 40),
]


@pytest.mark.parametrize("q,n,a_pad,kmax", SHAPES)
def test_every_tile_of_every_row_block_once(q, n, a_pad, kmax):
    plan, items = plan_and_items(q, n, a_pad, kmax)
    mb, nt, chunks = plan["m_blocks"], plan["n_tiles"], plan["chunks"]
    assert mb == -(-q // 128) and nt == -(-n // 256) and plan["k_blocks"] == a_pad // 64
    assert plan["grid"] % plan["cluster"] == 0 and plan["grid"] <= SMS
    seen = np.zeros((mb, nt), dtype=np.int32)
    lists = np.zeros((mb, chunks), dtype=np.int32)
    for cl, rank, m, chunk, tb, te, group, rows in items:
        assert 0 <= chunk < chunks and 0 <= tb <= te <= nt
        if m >= mb:
            continue  # phantom block of a partly filled cluster row: computes on zero fill, writes nothing
        seen[m, tb:te] += 1
        lists[m, chunk] += 1
    assert (seen == 1).all()
    assert (lists == 1).all()


@pytest.mark.parametrize("q,n,a_pad,kmax", SHAPES)
def test_chunks_tile_the_data_and_leave_room_for_k(q, n, a_pad, kmax):
    plan, items = plan_and_items(q, n, a_pad, kmax)
    by_chunk = {}
    for _, _, _, chunk, tb, te, _, _ in items:
        by_chunk.setdefault(int(chunk), set()).add((int(tb), int(te)))
    assert all(len(v) == 1 for v in by_chunk.values())          # a chunk has one tile range
    ranges = sorted(next(iter(v)) for v in by_chunk.values())
    assert ranges[0][0] == 0 and ranges[-1][1] == plan["n_tiles"]
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))  # contiguous, no overlap
    # every chunk keeps 32 keys per query; the certified survivor set is about 1.6 k
    if plan["n_tiles"] >= (kmax + 7) // 8:
        assert plan["chunks"] * plan["kp"] >= 2 * min(kmax, n) or plan["chunks"] == plan["n_tiles"]


@pytest.mark.parametrize("q,n,a_pad,kmax", SHAPES)
def test_realignment_groups_are_co_resident(q, n, a_pad, kmax):
    """Clusters that wait for each other on a (group, chunk) counter must be distinct clusters
    working on the group in the same position of their item lists, otherwise the wait of the
    pair kernel could never complete."""
    plan, items = plan_and_items(q, n, a_pad, kmax)
    if plan["cluster"] != 2:
        return
    nclusters = plan["grid"] // 2
    lead = items[items[:, 1] == 0]
    pos, nxt = {}, {}
    for cl, _, m, chunk, _, _, group, rows in lead:
        i = nxt.get(int(cl), 0)
        nxt[int(cl)] = i + 1
        pos.setdefault((int(group), int(chunk)), []).append((int(cl), i, int(rows)))
    whole_wave = plan["group_m"] // 2 * plan["chunks"] * 2 == plan["grid"]
    for (group, chunk), members in pos.items():
        cls = [c for c, _, _ in members]
        assert len(set(cls)) == len(cls) <= nclusters
        assert all(r == len(members) for _, _, r in members)      # the counter's target is the member count
        if whole_wave:
            assert len({i for _, i, _ in members}) == 1            # same wave for all of them


def test_wide_tile_only_for_long_rows(monkeypatch):
    lib = _lib()
    lib.dmlp_debug_topk_wide.argtypes = [C.c_int] * 5
    for a_pad, want in ((64, 0), (2048, 0), (4096, 1), (8192, 1)):
        assert lib.dmlp_debug_topk_wide(4096, 8192, a_pad, SMS, 16) == want
        plan, items = plan_and_items(4096, 8192, a_pad, 16)
        assert plan["cluster"] == 2 and len(items) > 0
    assert lib.dmlp_debug_topk_wide(100, 8192, 8192, SMS, 16) == 0   # single row block: one-CTA kernel
    monkeypatch.setenv("DMLP_TOPK_WIDE", "0")
    assert lib.dmlp_debug_topk_wide(4096, 8192, 8192, SMS, 16) == 0
    monkeypatch.setenv("DMLP_TOPK_WIDE", "1")
    assert lib.dmlp_debug_topk_wide(4096, 8192, 64, SMS, 16) == 1


def test_env_overrides_are_respected(monkeypatch):
    monkeypatch.setenv("DMLP_TOPK_CHUNKS", "4")
    monkeypatch.setenv("DMLP_TOPK_GMC", "18")
    plan, items = plan_and_items(65536, 65536, 8192, 16)
    assert plan["chunks"] == 4 and plan["group_m"] == 36 and plan["grid"] == 144
    monkeypatch.delenv("DMLP_TOPK_CHUNKS")
    monkeypatch.delenv("DMLP_TOPK_GMC")
    plan, _ = plan_and_items(65536, 65536, 8192, 16)
    assert plan["chunks"] == 2 and plan["group_m"] == 74 and plan["grid"] == 148


def test_random_shapes_cover_the_work_exactly_once():
    """300 random shapes (and SM counts: the planner is also used with fewer SMs, e.g. MIG slices):
    same invariants as above, including the co-residency of re-alignment groups."""
    rng = np.random.default_rng(2024)
    for _ in range(300):
        q = int(rng.choice([rng.integers(1, 300), rng.integers(300, 5000), rng.integers(5000, 70000)]))
        n = int(rng.choice([rng.integers(1, 600), rng.integers(600, 20000), rng.integers(20000, 300000)]))
        a_pad = 64 * int(rng.integers(1, 130))
        kmax = int(rng.choice([1, 5, 16, 40, 96]))
        sms = int(rng.choice([148, 148, 132, 64, 16, 2, 1]))
        plan, items = plan_and_items(q, n, a_pad, kmax, sms)
        mb, nt, chunks, cluster = plan["m_blocks"], plan["n_tiles"], plan["chunks"], plan["cluster"]
        assert 1 <= plan["grid"] <= max(sms, cluster) and plan["grid"] % cluster == 0, (q, n, a_pad, kmax, sms, plan)
        seen = np.zeros((mb, nt), dtype=np.int32)
        lists = np.zeros((mb, chunks), dtype=np.int32)
        groups = {}
        for cl, rank, m, chunk, tb, te, group, rows in items:
            if rank == 0:
                groups.setdefault((int(group), int(chunk)), []).append((int(cl), int(rows)))
            if m < mb:
                seen[m, tb:te] += 1
                lists[m, chunk] += 1
        assert (seen == 1).all() and (lists == 1).all(), (q, n, a_pad, kmax, sms, plan)
        for members in groups.values():
            cls = [c for c, _ in members]
            assert len(set(cls)) == len(cls) and all(r == len(members) for _, r in members), (q, n, a_pad, kmax, sms)


@pytest.mark.parametrize("q,n,a_pad,kmax,parts", [(65536, 65536, 8192, 1024, 4), (600, 40000, 256, 300, 2),
                                                  (600, 20000, 64, 300, 1), (600, 12000, 64, 128, 1),
                                                  (600, 70000, 512, 600, 3)])
def test_large_k_plans_split_the_data_into_parts(q, n, a_pad, kmax, parts):
    """k > 256: the filter runs as `parts` launches over row ranges of the data, every range with
    the same number of lists; in total at least one 32-key list per 8 of k."""
    lib = _lib()
    lib.dmlp_fused_supported.argtypes = [C.c_int, C.c_int64]
    assert lib.dmlp_fused_supported(kmax, n) == 1
    assert lib.dmlp_fused_supported(kmax, 24 * kmax) == 0 and lib.dmlp_fused_supported(96, 1) == 2
    assert lib.dmlp_fused_supported(1024, 32512) == 0 and lib.dmlp_fused_supported(1024, 32513) == 1
    arr = (C.c_int * 8)()
    assert lib.dmlp_topk_plan(q, n, a_pad, SMS, kmax, C.byref(arr)) == 0
    chunks = arr[3]
    assert chunks % parts == 0 and chunks * 8 >= kmax and chunks // parts <= 64
    assert arr[1] == -(-n // 256) and arr[5] <= SMS


def test_routing_rule_and_planner_agree():
    """Wherever dmlp_fused_supported sends a k to the filter path, the planner gives it at least one
    32-key list per 8 of k -- for any row length (one or two sets of epilogue warps, wide tile)."""
    lib = _lib()
    lib.dmlp_fused_supported.argtypes = [C.c_int, C.c_int64]
    arr = (C.c_int * 16)()
    checked = 0
    for kmax in (1, 8, 16, 96, 97, 128, 200, 256, 257, 300, 512, 513, 777, 1000, 1024):
        for n in (256, 4000, 8192, 16129, 20000, 32513, 70000, 300000):
            level = lib.dmlp_fused_supported(kmax, n)
            if level == 0:
                assert kmax > 96
                continue
            assert level == (2 if kmax <= 96 else 1)
            if kmax <= 96:
                continue          # (small k: the planner clamps to the tiles there are; the certificate flags what overflows)
            for a_pad in (64, 128, 1024, 8192):
                assert lib.dmlp_topk_plan_ex(2000, n, a_pad, SMS, kmax, C.byref(arr)) == 0
                lists, halves, parts, chunks = arr[3], arr[8], arr[9], arr[10]
                assert lists == chunks * halves * parts
                assert lists * 8 >= kmax, (kmax, n, a_pad, lists)
                assert chunks <= 32 and chunks * parts <= -(-n // 256)      # a list needs a tile of its own
                checked += 1
    assert checked > 100
    assert lib.dmlp_fused_supported(1025, 10 ** 9) == 0 and lib.dmlp_fused_supported(0, 100) == 0
