#!/usr/bin/env python3
"""Host -> device copy rate from pinned memory (what bounds the end-to-end step), with and without
binding the process to the GPU's NUMA node (utils/numa.py)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.utils import numa  # noqa: E402


def rate(gib=2, reps=5):
    host = torch.empty(gib << 30, dtype=torch.uint8).pin_memory()
    dev = torch.empty_like(host, device="cuda")
    best = 0.0
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dev.copy_(host, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, host.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    return best


props = torch.cuda.get_device_properties(0)
out = {"gpu": props.name, "cpus_allowed": len(os.sched_getaffinity(0)), "h2d_gbs_unbound": rate()}
os.environ["DMLP_NUMA_BIND"] = "1"
out["numa_node"] = numa.bind_to_gpu_node(0)
out["cpus_after_bind"] = len(os.sched_getaffinity(0))
out["h2d_gbs_bound"] = rate()   # new pinned buffer, allocated after the binding
print(json.dumps(out))
