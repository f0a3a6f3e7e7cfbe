"""Multi-GPU paths (need >= 2 GPUs): fused peer-memory engine and the NCCL+cuBLAS baseline
must both reproduce the CPU oracle.  Spawns one process per GPU with torch.distributed.run."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

WORKER = r'''
import os, sys, json
sys.path.insert(0, os.environ["DMLP_ROOT"])
import numpy as np, torch, torch.distributed as dist
from dmlp_b200.parallel import DistributedKNN
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
impl, n, q, a, kmin, kmax = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
grid = tuple(int(v) for v in sys.argv[7].split("x"))
eng = DistributedKNN(device=dev, impl=impl, grid=grid)
eng.synth_problem(n_total=n, q_total=q, a=a, lo=-100.0, hi=100.0, min_k=kmin, max_k=kmax, num_labels=7, seed=5)
for step in range(3):                      # several steps: epochs, flag reuse, buffer reuse
    checked = eng.verify_against_oracle(10 ** 9)
sampled = eng.verify_sampled(n_exact=24, n_cpu=3)   # the check bench.py runs at any size
assert sampled == 24 * eng.C + 3, sampled
assert eng.check_watchdog() == 0
ref_lab, ref_chk = eng.run_resident()
ref_chk = ref_chk.cpu().numpy().copy()
for wire in ("auto", "fp64"):
    eng.stage_host_inputs(wire=wire)
    for step in range(3):                  # buffers are overwritten every step: reader/writer flags
        lab, chk = eng.run_from_host()
        assert (chk.numpy() == ref_chk).all(), (wire, step)
assert eng.check_watchdog() == 0
print(json.dumps({"rank": rank, "checked": checked, "flagged": eng.last_flagged, "peer_fused": bool(getattr(eng, "peer_fused", False)), "fused_ok": bool(eng.fused_ok), "rows_fused": bool(eng.rows_fused)}))
eng.close()
dist.destroy_process_group()
'''


def _run(tmp_path, repo_root, world, *args):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, DMLP_ROOT=repo_root)
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", str(script), *map(str, args)],
                         capture_output=True, text=True, env=env, timeout=240)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    return res.stdout


def _world():
    n = torch.cuda.device_count()
    return 8 if n >= 8 else 4 if n >= 4 else 2 if n >= 2 else 0


S0, S1, S2 = (6000, 1000, 96, 1, 12), (3001, 515, 33, 2, 20), (20000, 700, 256, 16, 16)
CASES = [("fused", S0, "rows"), ("fused", S0, "cols"), ("fused", S1, "rows"), ("fused", S1, "cols"),
         ("fused", S2, "rows"), ("nccl_cublas", S0, "rows"), ("nccl_cublas", S1, "cols"),
         ("fused", S1, "2d"), ("nccl_cublas", S0, "2d")]


def _grid(kind, world):
    """rows: pure data sharding (R x 1); cols: pure query sharding (1 x C); 2d: needs >= 4 GPUs."""
    if kind == "rows":
        return world, 1
    if kind == "cols":
        return 1, world
    return (2, world // 2) if world >= 4 else None


@pytest.mark.parametrize("impl,shape,kind", CASES)
def test_distributed_matches_oracle(tmp_path, repo_root, impl, shape, kind):
    world = _world()
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    grid = _grid(kind, world)
    if grid is None:
        pytest.skip("a 2-D grid needs at least 4 GPUs")
    out = _run(tmp_path, repo_root, world, impl, *shape, f"{grid[0]}x{grid[1]}")
    assert out.count('"checked"') == world
    if impl == "fused" and grid[0] > 1:
        assert '"peer_fused": true' in out


def test_distributed_large_k(tmp_path, repo_root):
    """k up to 60 stays on the fused path (more chunks); k = 120 takes the exact kernels."""
    world = _world()
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    out = _run(tmp_path, repo_root, world, "fused", 20000, 300, 64, 30, 60, f"{world}x1")
    assert out.count('"checked"') == world and '"peer_fused": true' in out
    out = _run(tmp_path, repo_root, world, "fused", 2000, 300, 16, 100, 120, f"{world}x1")
    assert out.count('"checked"') == world and '"peer_fused": false' in out
    # 96 < k <= 1024 with queries sharded (1 x C): the filter path in several launches over the shard,
    # which the peer_gather kernel assembles first (the in-kernel gathers need a single launch)
    out = _run(tmp_path, repo_root, world, "fused", 40000, 600, 32, 200, 300, f"1x{world}")
    assert out.count('"checked"') == world and out.count('"fused_ok": true') == world
    assert '"rows_fused": true' not in out


def test_engine_binary_device_grids(tmp_path, repo_root):
    """The C++ driver (`engine`, one process for all GPUs) on R x C device grids vs its CPU backend.
    Where the box offers NVSwitch multicast, the R x 1 and 1 x C layouts signal through multicast
    flag pages (one multimem store per hand-shake, backend string "... nvls"); DMLP_NVLS=0 keeps
    the per-peer stores.  Both must print the CPU backend's lines."""
    import hashlib
    import json
    n_gpu = torch.cuda.device_count()
    if n_gpu < 2:
        pytest.skip("needs at least 2 GPUs")
    exe = os.path.join(repo_root, "engine")
    args = ["--json", "--synth", "30000", "900", "48", "-100", "100", "1", "20", "7", "5"]

    def run(env_extra):
        res = subprocess.run([exe, *args], capture_output=True, text=True, env=dict(os.environ, **env_extra), timeout=240)
        assert res.returncode == 0, res.stderr[-2000:]
        info = json.loads([ln for ln in res.stderr.splitlines() if ln.startswith("{")][-1])
        return hashlib.md5("".join(sorted(res.stdout.splitlines(True))).encode()).hexdigest(), info["backend"]

    want, _ = run({"DMLP_BACKEND": "cpu"})
    grids = [f"{n_gpu}x1", f"1x{n_gpu}"] + ([f"2x{n_gpu // 2}"] if n_gpu >= 4 else [])
    backends = []
    for g in grids:
        for nvls in ("1", "0"):
            got, backend = run({"DMLP_BACKEND": "gpu", "DMLP_GRID": g, "DMLP_NVLS": nvls, "DMLP_TRACE": "1"})
            assert got == want, (g, nvls, backend)
            assert nvls == "1" or "nvls" not in backend
            backends.append(backend)
    print("engine backends:", backends)
    assert run({"DMLP_BACKEND": "gpu_exact", "DMLP_GRID": f"{n_gpu}x1"})[0] == want
