#!/usr/bin/env python3
"""A/B of the filter kernel's item schedule: static waves (default) vs the dynamic scheduler
(DMLP_TOPK_DYN=1, atomic work counter) on the headline shape and on ragged shapes.

Each line: shape, configuration, plan, median ms of dist_topk (CUDA events, after warm-up).
The dynamic scheduler switches the re-alignment of co-resident clusters off, so the static
schedule is also timed with DMLP_TOPK_SYNC=0 to separate the two effects; "fine" doubles the
chunk count (items half as long: what a dynamic scheduler needs to even out a ragged tail)."""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.ops import cuda as K  # noqa: E402

SHAPES = [(65536, 65536, 8192), (70000, 65536, 8192), (70000, 100000, 1024), (70000, 200000, 64),
          (20000, 300000, 512)]
REPS = int(os.environ.get("AB_REPS", "5"))
KNOBS = ("DMLP_TOPK_DYN", "DMLP_TOPK_SYNC", "DMLP_TOPK_CHUNKS", "DMLP_TOPK_GMC")


def run(qh, xh, xn, env):
    for key in KNOBS:
        os.environ.pop(key, None)
    os.environ.update(env)
    ms = []
    for rep in range(REPS + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        K.dist_topk(qh, xh, xn, kmax=16)
        e1.record()
        torch.cuda.synchronize()
        if rep >= 2:
            ms.append(e0.elapsed_time(e1))
    return statistics.median(ms), min(ms)


for (Q, N, A) in SHAPES:
    x = K.synth_attrs(N, A, 0, -100.0, 100.0, 42, 0)
    qq = K.synth_attrs(Q, A, 0, -100.0, 100.0, 42, 2)
    st = K.new_stats()
    K.stats_update(x, st)
    K.stats_update(qq, st)
    xh, xn, _ = K.pack(x, st, st)
    qh, _, _ = K.pack(qq, st)
    del x, qq
    for key in KNOBS:
        os.environ.pop(key, None)
    plan = K.topk_plan(Q, N, K.pad64(A), kmax=16)
    fine = {"DMLP_TOPK_CHUNKS": str(min(2 * plan["chunks"], 32))}
    base = None
    for name, env in (("static", {}), ("static_nosync", {"DMLP_TOPK_SYNC": "0"}), ("dynamic", {"DMLP_TOPK_DYN": "1"}),
                      ("static_fine", dict(fine)), ("dynamic_fine", dict(fine, DMLP_TOPK_DYN="1"))):
        med, best = run(qh, xh, xn, env)
        for key in KNOBS:
            os.environ.pop(key, None)
        os.environ.update(env)
        pl = K.topk_plan(Q, N, K.pad64(A), kmax=16)
        base = med if base is None else base
        print(json.dumps({"shape": [Q, N, A], "config": name, "grid": pl["grid"], "group_m": pl["group_m"],
                          "chunks": pl["chunks"], "ms_median": round(med, 3), "ms_best": round(best, 3),
                          "vs_static": round(base / med, 4), "sync_timeouts": K.topk_sync_timeouts()}), flush=True)
    del xh, xn, qh
    torch.cuda.empty_cache()
