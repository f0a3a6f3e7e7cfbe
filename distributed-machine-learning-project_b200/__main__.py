"""`python -m dmlp_b200 < input.txt`: the reference driver's contract from Python.

Reads the reference's text grammar on stdin (common.cpp:93-117), runs the job, prints one
`Query <id> checksum: <h>` line per query on stdout (common.cpp:57-71) and
`Time taken: <ms> ms` on stderr (common.cpp:119-131).  --debug prints the label and the
neighbour list as the reference's -DDEBUG build does.
"""
import argparse
import sys
import time

import numpy as np

from .ops import cpu


def main(argv=None):
    p = argparse.ArgumentParser(prog="python -m dmlp_b200", description=__doc__.splitlines()[0])
    p.add_argument("--backend", choices=["auto", "cpu", "gpu"], default="auto")
    p.add_argument("--strategy", default="serial", help="CPU strategy: serial grid gather tree farm kdtree")
    p.add_argument("--ranks", type=int, default=1, help="CPU 'ranks' for the multi-rank strategies")
    p.add_argument("--threads", type=int, default=0)
    p.add_argument("--debug", action="store_true", help="engine.debug output: label and neighbour lists")
    args = p.parse_args(argv)

    data, labels, queries, k = cpu.parse_text(sys.stdin.buffer.read(), threads=args.threads)
    backend = args.backend
    if backend == "auto":
        try:
            import torch
            backend = "gpu" if torch.cuda.is_available() else "cpu"
        except Exception:
            backend = "cpu"

    t0 = time.perf_counter()
    if backend == "gpu":
        import torch
        from .models import knn
        res = knn(data, labels, queries, k, want_lists=args.debug)
        torch.cuda.synchronize()
        label, checksum = res.label.cpu().numpy(), res.checksums_u64()
        ids = res.ids.cpu().numpy() if args.debug else None
        dist = res.dist.cpu().numpy() if args.debug else None
        lists = None if not args.debug else [(ids[i, :max(int(k[i]), 0)], dist[i, :max(int(k[i]), 0)]) for i in range(len(k))]
    else:
        label, checksum, ids, dist, off = cpu.knn(data, labels, queries, k, strategy=args.strategy, ranks=args.ranks,
                                                  threads=args.threads)
        lists = [(ids[off[i]:off[i + 1]], dist[off[i]:off[i + 1]]) for i in range(len(k))] if args.debug else None
    ms = int((time.perf_counter() - t0) * 1e3)

    out = []
    for i in range(len(k)):
        if args.debug:
            out.append(f"Label for Query {i} : {int(label[i])}")
            out.append(f"Top-{max(int(k[i]), 0)} neighbors:")
            out.extend(f"{int(j)} : {float(d):g}" for j, d in zip(*lists[i]) if j >= 0)
        else:
            out.append(f"Query {i} checksum: {int(np.uint64(checksum[i]))}")
    sys.stdout.write("\n".join(out) + ("\n" if out else ""))
    print(f"Time taken: {ms} ms", file=sys.stderr)
    return 0


if __name__ == "__main__":
    sys.exit(main())
