#!/usr/bin/env python3
"""Pinned host -> device copy rate with every rank copying at the same time (what bounds the
end-to-end step on several GPUs of one host).  Launch with torch.distributed.run."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
host = torch.empty(2 << 30, dtype=torch.uint8).pin_memory()
dst = torch.empty_like(host, device=dev)
out = {}
for mode in ("alone", "together"):
    best = 0.0
    for rep in range(4):
        for r in range(world if mode == "alone" else 1):
            dist.barrier()
            if mode == "together" or r == rank:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dst.copy_(host, non_blocking=True)
                e1.record()
                torch.cuda.synchronize()
                if rep > 0:
                    best = max(best, host.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    t = torch.tensor([best], dtype=torch.float64, device=dev)
    allr = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allr, t)
    out[mode] = [round(float(v.item()), 1) for v in allr]
if rank == 0:
    print(json.dumps({"gpus": world, "h2d_gbs_alone_per_gpu": out["alone"], "h2d_gbs_all_at_once_per_gpu": out["together"],
                      "aggregate_all_at_once": round(sum(out["together"]), 1)}))
dist.destroy_process_group()
