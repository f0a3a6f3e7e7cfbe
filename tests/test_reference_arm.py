"""The reference arm of bench.py: the UNMODIFIED reference sources compiled against the MPI
stand-in (baseline/mpi_shim/mpi.h), on one rank and on several OS processes.

Skipped where the reference sources are not reachable (they are not part of this repository).
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench_reference  # noqa: E402
from dmlp_b200.ops import cpu  # noqa: E402


@pytest.fixture(scope="module")
def ref_bin():
    why = bench_reference.ensure_reference()
    if why:
        pytest.skip(why)
    return bench_reference.REF_BIN


def _problem_text(n, q, a, k, seed):
    data, labels, queries, kk = cpu.synth(n, q, a, -10.0, 10.0, k, k, 4, seed=seed)
    lines = [f"{n} {q} {a}"]
    lines += [" ".join([str(int(labels[i]))] + [f"{v:.6f}" for v in data[i]]) for i in range(n)]
    lines += [" ".join(["Q", str(int(kk[i]))] + [f"{v:.6f}" for v in queries[i]]) for i in range(q)]
    _, want_sum, *_ = cpu.knn(data, labels, queries, kk)
    return "\n".join(lines) + "\n", want_sum


def _run_ranks(ref_bin, world, text, tmp_path):
    sock = tmp_path / f"sock{world}"
    sock.mkdir()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), DMLP_MPI_DIR=str(sock))
        procs.append(subprocess.Popen([ref_bin], stdin=subprocess.PIPE if r == 0 else subprocess.DEVNULL,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env))
    outs = [None] * world
    outs[0] = procs[0].communicate(text, timeout=120)
    for r in range(1, world):
        outs[r] = procs[r].communicate(timeout=120)
    return [p.returncode for p in procs], outs


def _checksums(outs):
    got = {}
    for stdout, _ in outs:
        for line in stdout.splitlines():
            parts = line.split()          # "Query <id> checksum: <h>"
            if len(parts) == 4 and parts[0] == "Query":
                got.setdefault(int(parts[1]), set()).add(int(parts[3]))
    return got


@pytest.mark.parametrize("world", [1, 4, 9])
def test_reference_runs_on_square_grids_and_agrees_with_our_oracle(ref_bin, world, tmp_path):
    text, want = _problem_text(700, 36, 7, 5, seed=3)
    codes, outs = _run_ranks(ref_bin, world, text, tmp_path)
    assert codes == [0] * world, [o[1][-200:] for o in outs]
    assert "Time taken:" in outs[0][1]
    got = _checksums(outs)
    assert sorted(got) == list(range(36))
    assert all(len(v) == 1 for v in got.values())     # printed once per rank and iteration (B4), always the same
    assert np.array_equal(np.array([next(iter(got[i])) for i in range(36)], dtype=np.uint64), want)


def test_reference_aborts_on_a_non_square_grid(ref_bin, tmp_path):
    """on 2 ranks (2x1) the reference overruns its gather buffer.  The arm
    reports 'unavailable' for such sizes instead of timing undefined behaviour."""
    text, _ = _problem_text(700, 36, 7, 5, seed=3)
    codes, outs = _run_ranks(ref_bin, 2, text, tmp_path)
    assert codes[0] != 0 or _checksums(outs) == {}
    assert bench_reference._dims_create(2) == (2, 1) and bench_reference._dims_create(8) == (4, 2)
    assert bench_reference._dims_create(4) == (2, 2) and bench_reference._dims_create(1) == (1, 1)
    assert bench_reference._dims_create(6) == (3, 2) and bench_reference._dims_create(9) == (3, 3)


def test_bench_reference_arm_under_the_launcher(ref_bin):
    """bench.py --impl reference on 4 launched processes prints one JSON line with the metric."""
    env = dict(os.environ, DMLP_ROOT=ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "4", "--master-addr",
           "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "4",
           "--steps", "1", "--warmup", "0", "--num-data", "2000", "--num-attrs", "16", "--ref-queries", "5"]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 4 and d["value"] > 0 and d["config"]["queries_timed"] == 20
    cmd[cmd.index("--nproc-per-node") + 1] = "2"
    cmd[cmd.index("--gpus") + 1] = "2"
    cmd[cmd.index("--master-port") + 1] = "29632"
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0 and len(lines) == 1 and "unavailable" in json.loads(lines[0])
