"""Multi-process CPU run (gloo, world_size > 1) of the distributed data flow."""
import os
import subprocess
import sys

import pytest

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["DMLP_ROOT"])
import numpy as np, torch.distributed as dist
from dmlp_b200.ops import cpu
from dmlp_b200.parallel.cpu_engine import DistributedKNNCpu
dist.init_process_group("gloo")
R, C = (int(v) for v in sys.argv[1].split("x"))
n, q, a, kmax = 413, 57, 6, 9
data, labels, queries, k = cpu.synth(n, q, a, -5.0, 5.0, 1, kmax, 4, seed=12)   # same problem on every rank
eng = DistributedKNNCpu(grid=(R, C))
(pb, pe), _, (ob, oe), _ = eng.partition(n, q)
label, checksum = eng.run(data[pb:pe], labels[pb:pe], queries[ob:oe], k[ob:oe], n, q, kmax)
want_label, want_sum, *_ = cpu.knn(data, labels, queries, k)
assert (checksum == want_sum[ob:oe]).all() and (label == want_label[ob:oe]).all(), (dist.get_rank(), R, C)
owned = [None] * dist.get_world_size()
dist.all_gather_object(owned, (ob, oe))
assert sorted(i for b, e in owned for i in range(b, e)) == list(range(q))      # every query owned exactly once
sys.stdout.write(f"rank{dist.get_rank()}-ok\n")
sys.stdout.flush()
dist.destroy_process_group()
'''


@pytest.mark.parametrize("world,grid", [(2, "2x1"), (2, "1x2"), (4, "2x2"), (3, "3x1")])
def test_gloo_grid_matches_oracle(tmp_path, repo_root, world, grid):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, DMLP_ROOT=repo_root, OMP_NUM_THREADS="1")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1",
                          "--nnodes=1", f"--nproc-per-node={world}", str(script), grid],
                         capture_output=True, text=True, env=env, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert res.stdout.count("-ok") == world, res.stdout
