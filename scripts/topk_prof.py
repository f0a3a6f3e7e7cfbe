#!/usr/bin/env python3
"""Where the CTA-pair filter kernel waits (DMLP_TOPK_PROF counters, see include/dmlp_knn.h)."""
import argparse
import ctypes
import json
import os
import statistics
import sys

os.environ["DMLP_TOPK_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.ops import cuda as K  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--a", type=int, default=8192)
p.add_argument("--q", type=int, default=65536)
p.add_argument("--n", type=int, default=65536)
p.add_argument("--kmax", type=int, default=16)
args = p.parse_args()
x = K.synth_attrs(args.n, args.a, 0, -100.0, 100.0, 42, 0)
qq = K.synth_attrs(args.q, args.a, 0, -100.0, 100.0, 42, 2)
st, qst = K.new_stats(), K.new_stats()
K.stats_update(x, st)
K.stats_update(qq, st)
xh, xn, xe = K.pack(x, st, st)
qh, qn, qe = K.pack(qq, st, qst)
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K.dist_topk(qh, xh, xn, kmax=args.kmax)
    e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
buf = (ctypes.c_ulonglong * (256 * 16))()
lib = K._lib()
lib.dmlp_debug_topk_prof.argtypes = [ctypes.c_void_p, ctypes.c_int]
n = lib.dmlp_debug_topk_prof(buf, 256)
rows = [[buf[i * 16 + j] for j in range(16)] for i in range(n)]
lead = rows[0::2]
med = lambda v: statistics.median(v) if v else 0  # noqa: E731
out = {
    "ms": ms,
    "ctas": n,
    "mma_total_cycles": med([r[0] for r in lead]),
    "mma_wait_operands_frac": med([r[1] / max(r[0], 1) for r in lead]),
    "mma_wait_accumulator_frac": med([r[2] / max(r[0], 1) for r in lead]),
    "producer_wait_stage_frac": med([r[4] / max(r[3], 1) for r in lead]),
    "producer_wait_realign_frac": med([r[5] / max(r[3], 1) for r in lead]),
    "producer_wait_realign_frac_max": max([r[5] / max(r[3], 1) for r in lead], default=0),
    "epilogue_wait_frac": med([r[6] / max(r[7], 1) for r in rows]),
    "epilogue_stage_frac": med([r[8] / max(r[7], 1) for r in rows]),
    "epilogue_tmem_ld_frac": med([r[9] / max(r[7], 1) for r in rows]),
    "epilogue_score_frac": med([r[10] / max(r[7], 1) for r in rows]),
    "epilogue_compact_frac": med([r[11] / max(r[7], 1) for r in rows]),
    "epilogue_flush_frac": med([r[12] / max(r[7], 1) for r in rows]),
    "wide": os.environ.get("DMLP_TOPK_WIDE", "0"),
    "sync": os.environ.get("DMLP_TOPK_SYNC", "default"),
}
print(json.dumps(out))
