"""The `engine` binary keeps the reference driver contract."""
import os
import re
import subprocess
import sys

import pytest


@pytest.fixture(scope="module")
def engine_bins(repo_root):
    exe = os.path.join(repo_root, "engine")
    dbg = os.path.join(repo_root, "engine.debug")
    if not (os.path.exists(exe) and os.path.exists(dbg)):
        subprocess.check_call(["make", "-C", repo_root, "-j8", "all"], stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL)
    return exe, dbg


def _gen(tmp_path, repo_root, *extra):
    out = tmp_path / "in.txt"
    subprocess.check_call([sys.executable, os.path.join(repo_root, "tools", "generate_input.py"),
                           "--num_data", "12", "--num_queries", "3", "--num_attrs", "2", "--min", "0", "--max", "10",
                           "--num_labels", "3", "--seed", "42", "--output", str(out), *extra],
                          stdout=subprocess.DEVNULL)
    return out


def _run(exe, path, env_extra=None, args=()):
    env = dict(os.environ, DMLP_BACKEND="cpu")
    env.update(env_extra or {})
    with open(path, "rb") as f:
        return subprocess.run([exe, *args], stdin=f, capture_output=True, env=env, text=True)


def test_release_contract(engine_bins, tmp_path, repo_root):
    res = _run(engine_bins[0], _gen(tmp_path, repo_root, "--minK", "3", "--maxK", "3"))
    assert res.returncode == 0
    assert sorted(res.stdout.splitlines()) == sorted([
        "Query 0 checksum: 1048314657462637447",
        "Query 1 checksum: 17629266663427905930",
        "Query 2 checksum: 6436800073062518871"])
    assert re.fullmatch(r"Time taken: \d+ ms\n", res.stderr)


def test_debug_contract(engine_bins, tmp_path, repo_root):
    res = _run(engine_bins[1], _gen(tmp_path, repo_root, "--minK", "3", "--maxK", "3"))
    lines = res.stdout.splitlines()
    assert lines[0] == "Label for Query 0 : 0"
    assert lines[1] == "Top-3 neighbors:"
    assert lines[2].startswith("11 : 3.21078")


@pytest.mark.parametrize("env", [
    {"DMLP_STRATEGY": "grid", "DMLP_GRID": "2x1"}, {"DMLP_STRATEGY": "grid", "DMLP_GRID": "4x2"},
    {"DMLP_STRATEGY": "tree", "DMLP_RANKS": "8"}, {"DMLP_STRATEGY": "farm", "DMLP_RANKS": "3"},
    {"DMLP_STRATEGY": "gather", "DMLP_RANKS": "2"}, {"DMLP_STRATEGY": "kdtree"}])
def test_strategies_and_aos_api(engine_bins, tmp_path, repo_root, env):
    path = _gen(tmp_path, repo_root, "--minK", "2", "--maxK", "4")
    want = ["Query 0 checksum: 1698155651117006534", "Query 1 checksum: 17629266663427905930",
            "Query 2 checksum: 15745963786526292968"]
    for args in ((), ("--aos",)):
        res = _run(engine_bins[0], path, env, args)
        assert res.returncode == 0, res.stderr
        assert sorted(res.stdout.splitlines()) == sorted(want)  # each query exactly once


def test_bad_input_fails_cleanly(engine_bins, tmp_path):
    bad = tmp_path / "bad.in"
    bad.write_text("1 1 1\n0 1.0\nX 1 1.0\n")
    res = _run(engine_bins[0], bad)
    assert res.returncode != 0 and "wrongly formatted" in res.stderr


def test_run_bench_harness_flow(repo_root, tmp_path):
    """run_bench.sh keeps the reference CLI: usage on bad args, baseline cache, time compare, output diff."""
    sh = os.path.join(repo_root, "run_bench.sh")
    res = subprocess.run([sh], capture_output=True, text=True)
    assert res.returncode == 1 and "Usage" in res.stdout
    res = subprocess.run([sh, "7"], capture_output=True, text=True)
    assert res.returncode == 1 and "invalid CONFIG" in res.stdout
    for stale in ("outputs/test_1.err", "outputs/test_1.out", "inputs/input1.in"):
        p = os.path.join(repo_root, stale)
        if os.path.exists(p):
            os.remove(p)
    env = dict(os.environ, DMLP_BACKEND="cpu", DMLP_STRATEGY="gather", DMLP_RANKS="2",
               INPUT_ARGS_1="--num_data 300 --num_queries 40 --num_attrs 5 --min 0 --max 10 --minK 3 --maxK 3 --num_labels 3")
    res = subprocess.run([sh, "1"], capture_output=True, text=True, env=env)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "=== Performance Comparison ===" in res.stdout and "Results:        identical (40 queries)" in res.stdout
    res = subprocess.run([sh, "1"], capture_output=True, text=True, env=env)
    assert "Output found in cache. Skipping..." in res.stdout
    if os.path.exists(os.path.join(repo_root, "baseline", "_ref_build", "engine_ref")):
        # the baseline on four ranks of the MPI stand-in (its 2 x 2 grid): same results again
        for stale in ("outputs/test_1.err", "outputs/test_1.out"):
            os.remove(os.path.join(repo_root, stale))
        res = subprocess.run([sh, "1"], capture_output=True, text=True, env=dict(env, REF_RANKS="4"))
        assert res.returncode == 0 and "Results:        identical (40 queries)" in res.stdout, res.stdout + res.stderr
    for stale in ("outputs/test_1.err", "outputs/test_1.out", "inputs/input1.in"):
        os.remove(os.path.join(repo_root, stale))


def test_numa_binding_helpers(tmp_path, monkeypatch):
    """utils.numa: sysfs parsing on a fake tree; binding is best effort and never raises."""
    from dmlp_b200.utils import numa
    assert numa.parse_cpulist("0-3,8,10-11") == {0, 1, 2, 3, 8, 10, 11}
    assert numa.parse_cpulist("") == set() and numa.parse_cpulist("x-y") == set()
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices/system/node/node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("56-111\n")
    assert numa.gpu_numa_node("00000000:1B:00.0", str(tmp_path)) == 1
    assert numa.gpu_numa_node("0000:1b:00.0", str(tmp_path)) == 1
    assert numa.gpu_numa_node("0000:2c:00.0", str(tmp_path)) is None
    (dev / "numa_node").write_text("-1\n")
    assert numa.gpu_numa_node("0000:1b:00.0", str(tmp_path)) is None
    assert numa.node_cpus(1, str(tmp_path)) == set(range(56, 112))
    assert numa.bind_to_gpu_node(0) is None           # off unless DMLP_NUMA_BIND=1
    monkeypatch.setenv("DMLP_NUMA_BIND", "1")
    assert numa.bind_to_gpu_node(0) is None           # no GPU here: nothing changes, nothing raises


def test_python_module_cli_keeps_the_driver_contract(tmp_path, repo_root, engine_bins):
    """`python -m dmlp_b200 < input` prints what ./engine (and, with --debug, ./engine.debug) prints."""
    inp = tmp_path / "in.txt"
    subprocess.run([sys.executable, os.path.join(repo_root, "tools", "generate_input.py"), "--num_data", "400", "--num_queries", "30",
                    "--num_attrs", "6", "--min", "-5", "--max", "5", "--minK", "1", "--maxK", "7", "--num_labels", "3",
                    "--output", str(inp)], check=True, capture_output=True)
    env = dict(os.environ, PYTHONPATH=str(repo_root), DMLP_BACKEND="cpu")
    text = inp.read_bytes()
    for flag, binary in (([], "engine"), (["--debug"], "engine.debug")):
        py = subprocess.run([sys.executable, "-m", "dmlp_b200", "--backend", "cpu"] + flag, input=text,
                            capture_output=True, env=env, check=True)
        ref = subprocess.run([os.path.join(repo_root, binary)], input=text, capture_output=True, env=env, check=True)
        assert py.stdout == ref.stdout and len(py.stdout) > 0
        assert b"Time taken:" in py.stderr
    grid = subprocess.run([sys.executable, "-m", "dmlp_b200", "--backend", "cpu", "--strategy", "grid", "--ranks", "6"],
                          input=text, capture_output=True, env=env, check=True)
    assert grid.stdout == subprocess.run([os.path.join(repo_root, "engine")], input=text, capture_output=True, env=env).stdout
