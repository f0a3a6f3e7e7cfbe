#!/usr/bin/env python3
"""SASS mnemonic census of the dist_topk kernels (no GPU needed): proves tcgen05 / TMA / TMEM use.

usage: scripts/sass_census.py [lib] > profiles/sass_dist_topk.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "distributed-machine-learning-project_b200/lib/libdmlp_knn.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
print(f"SASS mnemonic census of the dist_topk kernels in {lib} (nvcc 12.9, sm_100a), cuobjdump -sass.\n"
      "UTCHMMA = tcgen05.mma, UTMALDG = TMA tile load, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit,\n"
      "SYNCS.* = mbarrier ops; LDG.E.128 / STG.E.128 in the pair kernel's gather warps are the peer-memory\n"
      "(NVLink) loads and local stores of the fused all-gathers (K4, data rows); ATOMG/RED = row-block and\n"
      "tile counters.  dist_topk_pair_kernel<true> is the wide (256 x 512 per pair) variant.\n")
name, counts = None, None


def flush():
    if name and "dist_topk" in name:
        print(name)
        print(f"  instructions: {sum(counts.values())}")
        keep = re.compile(r"^(UTC|UTMA|LDTM|SYNCS|ATOM|RED|LDG|STG|LDS|STS|BAR|MEMBAR|ERRBAR|FENCE|ELECT|VOTE|CALL|CCTL|UBLKCP)")
        for k in sorted(counts):
            if keep.match(k):
                print(f"    {k:32s} {counts[k]}")
        print()


for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        name, counts = m.group(1), collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Za-z0-9_.]*)", line)
    if m and counts is not None:
        counts[m.group(1)] += 1
flush()
