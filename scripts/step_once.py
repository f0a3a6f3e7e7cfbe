#!/usr/bin/env python3
"""Two resident steps of the headline job on one GPU (for ncu: skip the first step's launches).

  python scripts/step_once.py [--a 8192 --q 65536 --n 65536] [--steps 2]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.parallel import DistributedKNN  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--a", type=int, default=8192)
p.add_argument("--q", type=int, default=65536)
p.add_argument("--n", type=int, default=65536)
p.add_argument("--kmin", type=int, default=4)
p.add_argument("--kmax", type=int, default=16)
p.add_argument("--steps", type=int, default=2)
args = p.parse_args()
eng = DistributedKNN(device=torch.device("cuda", 0))
eng.synth_problem(n_total=args.n, q_total=args.q, a=args.a, lo=-100.0, hi=100.0, min_k=args.kmin, max_k=args.kmax,
                  num_labels=10, seed=42)
torch.cuda.synchronize()
for _ in range(args.steps):
    eng.run_resident()
torch.cuda.synchronize()
print("flagged", eng.ensure_repaired(), "launches", eng.launch_count)
