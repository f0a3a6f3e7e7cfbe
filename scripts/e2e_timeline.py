#!/usr/bin/env python3
"""Timeline of the pipelined end-to-end loop on one GPU: when each step's copies and kernels
start and end (CUDA events against a common origin)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.parallel import DistributedKNN  # noqa: E402

eng = DistributedKNN(device=torch.device("cuda", 0))
eng.synth_problem(n_total=65536, q_total=65536, a=8192, lo=-100.0, hi=100.0, min_k=4, max_k=16, num_labels=10, seed=42)
eng.stage_host_inputs()
for _ in range(2):
    eng.run_from_host()
torch.cuda.synchronize()
origin = torch.cuda.Event(enable_timing=True)
origin.record()
rows = []
orig_copy = eng._copy_pieces


def traced_copy(pieces, slot):
    e0 = torch.cuda.Event(enable_timing=True)
    e0.record(eng.copy_stream)
    ev = orig_copy(pieces, slot)
    e1 = torch.cuda.Event(enable_timing=True)
    e1.record(eng.copy_stream)
    rows.append(("copy", slot, e0, e1))
    return ev


eng._copy_pieces = traced_copy
steps = 6
for i in range(steps):
    h0 = time.perf_counter()
    c0 = torch.cuda.Event(enable_timing=True)
    c0.record()
    eng.run_from_host(prefetch_next=i + 1 < steps)
    c1 = torch.cuda.Event(enable_timing=True)
    c1.record()
    rows.append(("step", i, c0, c1, (time.perf_counter() - h0) * 1e3))
torch.cuda.synchronize()
for r in rows:
    print(json.dumps({"what": r[0], "id": r[1], "begin_ms": round(origin.elapsed_time(r[2]), 2),
                      "end_ms": round(origin.elapsed_time(r[3]), 2), **({"host_ms": round(r[4], 2)} if len(r) > 4 else {})}))
