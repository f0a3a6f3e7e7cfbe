"""Golden vectors and property tests for the CPU backends."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

from dmlp_b200.ops import cpu

STRATEGIES = ["serial", "grid", "gather", "tree", "farm", "kdtree"]


def gen_reference_stream(tmp_path, repo_root, **kw):
    out = tmp_path / "in.txt"
    args = [sys.executable, os.path.join(repo_root, "tools", "generate_input.py")]
    for key, val in kw.items():
        args += [f"--{key}", str(val)]
    args += ["--output", str(out)]
    subprocess.check_call(args, stdout=subprocess.DEVNULL)
    return out.read_bytes()


def brute_numpy(data, labels, queries, k):
    """Independent restatement of the spec in plain Python/numpy."""
    sums = []
    for qi in range(queries.shape[0]):
        acc = np.zeros(data.shape[0])
        for j in range(data.shape[1]):  # sequential, non-fused: same rounding as the reference
            d = queries[qi, j] - data[:, j]
            acc = acc + d * d
        order = sorted(range(data.shape[0]), key=lambda r: (acc[r], -r))[: max(int(k[qi]), 0)]
        votes = {}
        for r in order:
            votes[int(labels[r])] = votes.get(int(labels[r]), 0) + 1
        label = -1
        best = 0
        for lab, cnt in votes.items():
            if cnt > best or (cnt == best and lab > label):
                best, label = cnt, lab
        h = 1469598103934665603
        h = ((h ^ (label & (2**64 - 1))) * 1099511628211) % 2**64
        for r in order:
            h = ((h ^ (r + 1)) * 1099511628211) % 2**64
        sums.append(h)
    return np.array(sums, dtype=np.uint64)


def test_golden_constant_k(tmp_path, repo_root):
    text = gen_reference_stream(tmp_path, repo_root, num_data=12, num_queries=3, num_attrs=2, min=0, max=10,
                                minK=3, maxK=3, num_labels=3, seed=42)
    data, labels, queries, k = cpu.parse_text(text)
    label, checksum, ids, dist, off = cpu.knn(data, labels, queries, k)
    assert checksum.tolist() == [1048314657462637447, 17629266663427905930, 6436800073062518871]
    assert label.tolist() == [0, 0, 2]
    assert ids[:3].tolist() == [11, 9, 1]
    np.testing.assert_allclose(dist[:3], [3.21078, 14.6927, 15.3981], rtol=1e-5)


def test_golden_variable_k(tmp_path, repo_root):
    text = gen_reference_stream(tmp_path, repo_root, num_data=12, num_queries=3, num_attrs=2, min=0, max=10,
                                minK=2, maxK=4, num_labels=3, seed=42)
    data, labels, queries, k = cpu.parse_text(text)
    for s in STRATEGIES:
        _, checksum, *_ = cpu.knn(data, labels, queries, k, strategy=s, ranks=3)
        assert checksum.tolist() == [1698155651117006534, 17629266663427905930, 15745963786526292968], s


@pytest.mark.parametrize("strategy,ranks,rows,cols", [
    ("serial", 1, 0, 0), ("grid", 2, 0, 0), ("grid", 8, 0, 0), ("grid", 8, 8, 1), ("grid", 8, 1, 8),
    ("grid", 6, 3, 2), ("grid", 6, 2, 3), ("gather", 5, 0, 0), ("tree", 7, 0, 0), ("tree", 8, 0, 0),
    ("farm", 4, 0, 0), ("kdtree", 1, 0, 0)])
def test_strategies_agree_with_spec(strategy, ranks, rows, cols):
    data, labels, queries, k = cpu.synth(157, 23, 5, -3.0, 3.0, 1, 9, 4, seed=7)
    want = brute_numpy(data, labels, queries, k)
    _, got, *_ = cpu.knn(data, labels, queries, k, strategy=strategy, ranks=ranks, rows=rows, cols=cols)
    assert got.tolist() == want.tolist()


def test_ties_prefer_larger_id_and_label():
    # four identical points: ties resolved by larger id; label vote ties -> larger label
    data = np.zeros((4, 3))
    labels = np.array([5, 1, 9, 1], dtype=np.int32)
    queries = np.ones((1, 3))
    for s in STRATEGIES:
        label, _, ids, *_ = cpu.knn(data, labels, queries, np.array([2], np.int32), strategy=s, ranks=2)
        assert ids.tolist() == [3, 2], s
        assert label.tolist() == [9], s  # labels {1, 9} tie 1-1 -> 9


def test_k_equals_num_data_and_shard_smaller_than_k():
    data, labels, queries, k = cpu.synth(9, 4, 3, 0.0, 1.0, 9, 9, 3, seed=3)
    want = brute_numpy(data, labels, queries, k)
    for s, ranks in [("grid", 4), ("gather", 8), ("tree", 8), ("farm", 3), ("kdtree", 1)]:
        _, got, ids, *_ = cpu.knn(data, labels, queries, k, strategy=s, ranks=ranks)
        assert got.tolist() == want.tolist(), s
        assert (ids >= 0).all()


def test_edge_shapes():
    data, labels, queries, k = cpu.synth(10, 0, 2, 0.0, 1.0, 1, 3, 2, seed=1)
    label, checksum, ids, dist, off = cpu.knn(data, labels, queries, k, strategy="grid", ranks=4)
    assert label.shape == (0,) and off.tolist() == [0]
    data, labels, queries, k = cpu.synth(10, 3, 1, 0.0, 1.0, 1, 3, 1, seed=1)
    want = brute_numpy(data, labels, queries, k)
    assert cpu.knn(data, labels, queries, k, strategy="tree", ranks=3)[1].tolist() == want.tolist()


def test_merge_operator_is_commutative_and_associative():
    rng = np.random.default_rng(0)
    k = 6

    def rand_list(base):
        d = np.sort(rng.integers(0, 5, k).astype(np.float64))
        i = base + np.arange(k, dtype=np.int32)
        order = sorted(range(k), key=lambda t: (d[t], -i[t]))
        return d[order], i[order].astype(np.int32)
    a, b, c = rand_list(0), rand_list(100), rand_list(200)
    ab = cpu.merge_topk(*a, *b)
    ba = cpu.merge_topk(*b, *a)
    assert ab[1].tolist() == ba[1].tolist()
    left = cpu.merge_topk(*cpu.merge_topk(*a, *b), *c)
    right = cpu.merge_topk(*a, *cpu.merge_topk(*b, *c))
    assert left[1].tolist() == right[1].tolist()


def test_parser_errors_match_reference_messages():
    with pytest.raises(RuntimeError, match="Line is wrongly formatted"):
        cpu.parse_text("1 1 1\n0 1.0\n2 1.0\n")
    with pytest.raises(RuntimeError, match="Line is empty"):
        cpu.parse_text("2 0 1\n0 1.0\n\n")
    # 'Q4 ...' and 'Q 4 ...' both parse (reference strips the leading Q only)
    _, _, _, k = cpu.parse_text("1 2 1\n0 1.0\nQ1 0.5\nQ 1 0.5\n")
    assert k.tolist() == [1, 1]


def test_parse_update_record():
    # reference: parse_update (common.cpp:46-55): an id, then as many values as the line holds
    ident, vals = cpu.parse_update("17 1.5 -2.25 3e2 +4\n")
    assert ident == 17 and vals.tolist() == [1.5, -2.25, 300.0, 4.0]
    ident, vals = cpu.parse_update("3")
    assert ident == 3 and vals.size == 0
    with pytest.raises(RuntimeError, match="Line is empty"):
        cpu.parse_update("   \n")


def test_exact_distance_is_sequential_non_fused():
    rng = np.random.default_rng(1)
    a, b = rng.uniform(-100, 100, 257), rng.uniform(-100, 100, 257)
    acc = 0.0
    for x, y in zip(a, b):
        acc = acc + (x - y) * (x - y)
    assert cpu.exact_sqdist(a, b) == acc


def test_synth_matches_text_round_trip(tmp_path, repo_root):
    data, labels, queries, k = cpu.synth(20, 5, 3, -100.0, 100.0, 1, 4, 3, seed=11)
    out = tmp_path / "hash.in"
    subprocess.check_call([sys.executable, os.path.join(repo_root, "tools", "generate_input.py"),
                           "--num_data", "20", "--num_queries", "5", "--num_attrs", "3", "--min", "-100",
                           "--max", "100", "--minK", "1", "--maxK", "4", "--num_labels", "3", "--seed", "11",
                           "--stream", "hash", "--output", str(out)], stdout=subprocess.DEVNULL)
    d2, l2, q2, k2 = cpu.parse_text(out.read_bytes())
    assert (d2 == data).all() and (q2 == queries).all() and (l2 == labels).all() and (k2 == k).all()


@pytest.mark.slow
def test_midsize_md5(tmp_path, repo_root):
    text = gen_reference_stream(tmp_path, repo_root, num_data=100000, num_queries=1000, num_attrs=64, min=-100,
                                max=100, minK=16, maxK=16, num_labels=10, seed=42)
    data, labels, queries, k = cpu.parse_text(text)
    _, checksum, *_ = cpu.knn(data, labels, queries, k)
    lines = sorted(f"Query {i} checksum: {int(c)}\n" for i, c in enumerate(checksum))
    assert hashlib.md5("".join(lines).encode()).hexdigest() == "8cf1102ad3eb629255188098e0cfb4a1"


def _numpy_lists(data, queries, k, kmax, id_base):
    """Independent restatement of the spec: sequential fp64 sum per pair (vectorised over the
    points only), order (distance ascending, larger id first), sentinel padding."""
    q, n = queries.shape[0], data.shape[0]
    dist = np.full((q, kmax), 1e18)
    ids = np.full((q, kmax), -1, np.int32)
    for i in range(q):
        acc = np.zeros(n)
        for c in range(data.shape[1]):
            d = queries[i, c] - data[:, c]
            acc = acc + d * d
        order = np.lexsort((-np.arange(n), acc))[:k[i]]
        dist[i, :len(order)] = acc[order]
        ids[i, :len(order)] = order + id_base
    return dist, ids


@pytest.mark.parametrize("n,q,a,kmax,seed", [(37, 5, 3, 6, 0), (64, 16, 8, 4, 1), (203, 21, 17, 9, 2), (8, 1, 1, 8, 3),
                                             (7, 19, 5, 3, 4), (1000, 33, 2, 12, 5), (129, 4, 64, 1, 6)])
def test_blocked_simd_kernel_matches_a_numpy_restatement(n, q, a, kmax, seed):
    """local_topk_many (eight points per SIMD register, 16 queries per pass, row counts that
    are not multiples of either) against a numpy restatement; duplicated points force ties."""
    rng = np.random.default_rng(seed)
    data = np.round(rng.uniform(-3, 3, (n, a)), 2)
    data[n // 2:] = data[:n - n // 2]          # duplicates: equal distances, the larger id must win
    labels = rng.integers(0, 4, n).astype(np.int32)
    queries = np.round(rng.uniform(-3, 3, (q, a)), 2)
    k = rng.integers(1, kmax + 1, q).astype(np.int32)
    k = np.minimum(k, n)
    want_d, want_i = _numpy_lists(data, queries, k, kmax, 11)
    for threads in (1, 3):
        od, oi, ol = cpu.local_lists(data, labels, 11, queries, k, kmax, threads=threads)
        assert np.array_equal(oi, want_i)
        assert np.array_equal(od, want_d)      # bit for bit
        real = oi >= 0
        assert np.array_equal(ol[real], labels[oi[real] - 11])


def test_copy_order_planner_beats_plain_alternation():
    """End-to-end schedule: the simulated makespan of the chosen interleaving of query chunks and
    data parts is never worse than plain alternation, and every piece appears exactly once."""
    from dmlp_b200.parallel.engine import best_copy_order, simulate_copy_order
    q, x = [9472] * 6 + [65536 - 6 * 9472], [8192] * 8
    order = best_copy_order(q, x, 8192)
    assert sorted(order) == sorted([("q", i) for i in range(7)] + [("x", i) for i in range(8)])
    assert [i for k, i in order if k == "q"] == list(range(7)) and [i for k, i in order if k == "x"] == list(range(8))
    alt = [("q", 0), ("x", 0)] + [p for i in range(1, 8) for p in ([("x", i)] + ([("q", i)] if i < 7 else []))]
    assert simulate_copy_order(order, q, x, 8192) <= simulate_copy_order(alt, q, x, 8192)
    # too many interleavings to enumerate: falls back to alternation, still a valid order
    big = best_copy_order([100] * 20, [100] * 20, 64)
    assert len(big) == 40 and big[0] == ("q", 0) and big[1] == ("x", 0)


def test_decimal6_correction_step_equals_division():
    """value = i / 1e6: one multiply by RN(1e-6) plus the exact-remainder correction step gives the
    correctly rounded quotient (what the kernels compute with three fp64 operations).  All 2^32
    integers were checked once in C++; here a dense sample plus the extremes, with numpy's
    float64 (math.fma is not available before Python 3.13, so the remainder is formed exactly
    with integers)."""
    from fractions import Fraction
    rng = np.random.default_rng(5)
    ints = np.concatenate([rng.integers(-2**31, 2**31, 20000), np.arange(-1000, 1000),
                           np.array([-2**31, 2**31 - 1, 10**8, -10**8, 99999999, 123456789])])
    y = 1e-6
    for i in ints.tolist():
        v0 = float(i) * y
        # r = fma(-1e6, v0, i) is exact: compute it in rationals and check it is a double
        r_exact = Fraction(i) - Fraction(10**6) * Fraction(v0)
        r = float(r_exact)
        assert Fraction(r) == r_exact
        # v = fma(r, y, v0): one rounding of the exact sum
        v_exact = Fraction(r) * Fraction(y) + Fraction(v0)
        v = float(v_exact)          # Fraction -> float rounds to nearest even
        assert v == i / 1e6, i


def test_host_decimal6_encoding():
    """The loader's integer form (Problem::encode_decimal6, used by `./engine` to halve its PCIe
    bytes): exact for generator / text-grammar values, refused for anything else."""
    data, _, queries, _ = cpu.synth(3000, 50, 37, -100.0, 100.0, 1, 1, 2, seed=3)
    enc = cpu.encode_decimal6(data)
    assert enc is not None and enc.dtype == np.int32
    assert (enc.astype(np.float64) / 1e6 == data).all()
    assert (enc == np.rint(data * 1e6).astype(np.int64)).all()
    big = np.random.default_rng(1).integers(-2147483647, 2147483647, size=200000).astype(np.float64) / 1e6
    enc = cpu.encode_decimal6(big, threads=3)                 # threaded path, the whole int32 range
    assert enc is not None and (enc.astype(np.float64) / 1e6 == big).all()
    assert cpu.encode_decimal6(np.array([0.0, -0.0, 1.5])) is not None     # -0.0 travels as 0
    assert cpu.encode_decimal6(np.array([0.1234567])) is None               # seven decimals
    assert cpu.encode_decimal6(np.array([1.0, 5000.0])) is None             # 5e9 does not fit
    assert cpu.encode_decimal6(np.array([np.nan])) is None and cpu.encode_decimal6(np.array([np.inf])) is None


@pytest.mark.parametrize("a,bits", [(64, 28), (33, 28), (5, 9), (100, 31), (7, 2), (64, 32), (1, 17)])
def test_host_bit_packing_matches_the_wire_format(a, bits):
    """Host twin of the device's pack_bits / unpack_bits: value i travels as the `bits`-bit number
    i + 2^(bits-1), rows are little-endian bit streams padded to whole 32-bit words."""
    rng = np.random.default_rng(a * 100 + bits)
    half = 1 << (bits - 1)
    vals = rng.integers(-half, half, size=(57, a), dtype=np.int64)
    vals[0, :] = -half
    vals[1, :] = half - 1
    v32 = vals.astype(np.int32)
    words = cpu.pack_bits(v32, bits)
    assert words.shape == (57, (a * bits + 31) // 32)
    u = (vals + half).astype(object)
    for r in (0, 1, 2, 30, 56):
        stream = 0
        for c in range(a):
            stream |= int(u[r, c]) << (c * bits)
        expect = [(stream >> (32 * w)) & 0xFFFFFFFF for w in range(words.shape[1])]
        assert words[r].tolist() == expect, r
    assert (cpu.unpack_bits(words, a, bits) == v32).all()
    if bits < 32:    # the width rule is symmetric: bit_length(max |i|) + 1 (so -2^(bits-1) itself asks for one more)
        assert cpu.decimal6_bits(np.maximum(v32, -(half - 1))) == bits
        assert cpu.decimal6_bits(v32) == bits + 1
    data, *_ = cpu.synth(200, 1, 16, -100.0, 100.0, 1, 1, 2, seed=5)
    assert cpu.decimal6_bits(cpu.encode_decimal6(data)) == 28   # [-100, 100] with six decimals
