#!/usr/bin/env python3
"""Large k (96 < k <= 1024): the filter + certified re-rank path against the direct fp64 kernel.

One JSON line per shape: device time (CUDA events, second call) of K.knn_local and of K.exact on
the same resident operands, flagged queries, lists per query; the direct kernel is timed on the
first `--exact-q` queries and scaled (it is linear in the number of queries)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.ops import cuda as K  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--exact-q", type=int, default=1024)
args = p.parse_args()

SHAPES = [(65536, 65536, 8192, 300), (65536, 65536, 8192, 1000), (16384, 262144, 1024, 300),
          (16384, 262144, 1024, 1000), (16384, 1000000, 64, 1000), (65536, 65536, 8192, 96)]


def timed(fn):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), out


for (Q, N, A, kmax) in SHAPES:
    x = K.synth_attrs(N, A, 0, -100.0, 100.0, 42, 0)
    lab = K.synth_labels(N, 0, 10, 42)
    qq = K.synth_attrs(Q, A, 0, -100.0, 100.0, 42, 2)
    k = K.synth_k(Q, 0, max(1, kmax // 2), kmax, 42)
    plan = K.topk_plan(Q, N, K.pad64(A), kmax=kmax)
    ms_f, out = timed(lambda: K.knn_local(qq, k, x, lab, kmax))
    nflag = int(out[4].item())
    qe = min(args.exact_q, Q)
    ms_e, out_e = timed(lambda: K.exact(qq[:qe], x, lab, k[:qe], kmax))
    same = bool(torch.equal(out[1][:qe], out_e[1])) and bool(torch.equal(out[0][:qe], out_e[0]))
    print(json.dumps({"shape": [Q, N, A], "kmax": kmax, "lists_per_query": plan["chunks"],
                      "fused_ms": round(ms_f, 2), "flagged": nflag,
                      "exact_ms_scaled": round(ms_e * Q / qe, 1), "exact_ms_measured": round(ms_e, 2), "exact_q": qe,
                      "speedup": round(ms_e * Q / qe / ms_f, 1), "first_rows_identical": same}), flush=True)
    del x, lab, qq, k, out, out_e
    torch.cuda.empty_cache()
