#!/usr/bin/env python3
"""Per-phase device times of the single-GPU pipeline (CUDA events, median of reps)."""
import argparse
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dmlp_b200.ops import cuda as K  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--a", type=int, default=8192)
p.add_argument("--q", type=int, default=65536)
p.add_argument("--n", type=int, default=65536)
p.add_argument("--kmin", type=int, default=4)
p.add_argument("--kmax", type=int, default=16)
p.add_argument("--reps", type=int, default=5)
p.add_argument("--dec6", action="store_true", help="data rows resident in decimal6 form (int32)")
args = p.parse_args()
A, Q, N = args.a, args.q, args.n
x = K.synth_attrs(N, A, 0, -100.0, 100.0, 42, 0)
lab = K.synth_labels(N, 0, 10, 42)
qq = K.synth_attrs(Q, A, 0, -100.0, 100.0, 42, 2)
k = K.synth_k(Q, 0, args.kmin, args.kmax, 42)
st, qst = K.new_stats(), K.new_stats()
xi = K.encode_decimal6(x)[0] if args.dec6 else None
flags_t, nflag_t = torch.empty(Q, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
lists_t = (torch.empty((Q, args.kmax), dtype=torch.float64, device="cuda"), torch.empty((Q, args.kmax), dtype=torch.int32, device="cuda"), torch.empty((Q, args.kmax), dtype=torch.int32, device="cuda"))
times = {}


def timed(name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    times.setdefault(name, []).append(e0.elapsed_time(e1))
    return out


for rep in range(args.reps + 2):
    if rep == 2:
        times.clear()
    K.stats_reset(st)
    K.stats_reset(qst)
    timed("stats", lambda: (K.stats_update(x, st), K.stats_update(qq, st)))
    xh, xn, xe = timed("pack_x", lambda: K.pack_i32(xi, st, st) if args.dec6 else K.pack(x, st, st))
    qh, qn, qe = timed("pack_q", lambda: K.pack(qq, st, qst))
    cand = timed("dist_topk", lambda: K.dist_topk(qh, xh, xn, kmax=args.kmax))
    if args.dec6:
        nflag_t.zero_()
        timed("rerank", lambda: K.rerank_ex(cand, qq, None, lab, k, qn, qe, st, st, 0, args.kmax, flags_t, nflag_t,
                                            out=lists_t, x_i32=xi))
        od, oi, ol, flags, nflag = lists_t + (flags_t, nflag_t)
    else:
        od, oi, ol, flags, nflag = timed("rerank", lambda: K.rerank(cand, qq, x, lab, k, qn, qe, st, st, 0, args.kmax))
    timed("finalize", lambda: K.finalize(od[None], oi[None], ol[None], k))
import ctypes
wbuf = torch.zeros(Q, dtype=torch.int32, device="cuda")
K._lib().dmlp_debug_rerank_w.argtypes = [ctypes.c_void_p]
K._lib().dmlp_debug_rerank_w(ctypes.c_void_p(wbuf.data_ptr()))
K.rerank(cand, qq, x, lab, k, qn, qe, st, st, 0, args.kmax)
torch.cuda.synchronize()
K._lib().dmlp_debug_rerank_w(None)
wf = wbuf.float()
wstats = {"mean": wf.mean().item(), "p50": wf.median().item(), "p99": wf.kthvalue(int(0.99 * Q)).values.item(),
          "max": wf.max().item(), "k_mean": k.float().mean().item()}
plan = K.topk_plan(Q, N, xh.shape[1], kmax=args.kmax)
med = {n: statistics.median(v) for n, v in times.items()}
flops = 2.0 * Q * N * K.pad64(A)
print(json.dumps({"shape": [Q, N, A], "plan": plan, "ms": med, "total_ms": sum(med.values()),
                  "dist_topk_tflops": flops / (med["dist_topk"] * 1e-3) / 1e12, "flagged": int(nflag.item()), "survivors": wstats}))
