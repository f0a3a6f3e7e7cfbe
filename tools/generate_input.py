#!/usr/bin/env python3
"""Input generator with the reference tool's command line and file format.

Same ten options, same validation and the same text grammar as the reference
generator (reference: generate_input.py:26-48, 11-21): header ``N Q A``, then
``label a0 .. aN`` per datapoint and ``Q k a0 .. aN`` per query with six
decimals.  For a given ``--seed`` the default (``--stream python``) reproduces
the reference's output byte for byte because it draws from ``random`` in the
same order; it streams lines to disk instead of building the file in memory.

``--stream hash`` uses the engine's counter-based generator instead (the one the
device fill kernel implements, include/dmlp/io.h), which is what the benchmarks
use for problems too large for a text round trip.  With ``--native`` the hashed
stream is written by the C++ tool ``tools/dmlp-gen`` when it has been built.
"""
import argparse
import os
import random
import subprocess
import sys

MASK = (1 << 64) - 1


def _hash(seed, stream, index):
    z = (seed * 0x9E3779B97F4A7C15 + stream * 0xD1B54A32D192ED03 + index) & MASK
    z = (z + 0x9E3779B97F4A7C15) & MASK
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return z ^ (z >> 31)


def _attr(h, lo, hi):
    u = (h >> 11) * (1.0 / 9007199254740992.0)
    x = lo + (hi - lo) * u
    s = x * 1e6
    i = int(s - 0.5) if s < 0 else int(s + 0.5)
    return i / 1e6


def write_python_stream(f, a):
    f.write(f"{a.num_data} {a.num_queries} {a.num_attrs}\n")
    uni, rint = random.uniform, random.randint
    for _ in range(a.num_data):
        label = rint(0, a.num_labels - 1)
        f.write(str(label) + " " + " ".join("%.6f" % uni(a.min, a.max) for _ in range(a.num_attrs)) + "\n")
    kmax = min(a.maxK, a.num_data)
    for _ in range(a.num_queries):
        k = rint(a.minK, kmax)
        f.write("Q " + str(k) + " " + " ".join("%.6f" % uni(a.min, a.max) for _ in range(a.num_attrs)) + "\n")


def write_hash_stream(f, a):
    f.write(f"{a.num_data} {a.num_queries} {a.num_attrs}\n")
    A = a.num_attrs
    for i in range(a.num_data):
        label = _hash(a.seed, 1, i) % a.num_labels
        f.write(str(label) + " " + " ".join("%.6f" % _attr(_hash(a.seed, 0, i * A + j), a.min, a.max) for j in range(A)) + "\n")
    kmax = min(a.maxK, a.num_data)
    kmin = max(min(a.minK, kmax), 0)
    for i in range(a.num_queries):
        k = kmin + _hash(a.seed, 3, i) % (kmax - kmin + 1)
        f.write("Q " + str(k) + " " + " ".join("%.6f" % _attr(_hash(a.seed, 2, i * A + j), a.min, a.max) for j in range(A)) + "\n")


def main():
    p = argparse.ArgumentParser(description="Generate input for the k-NN engine.")
    p.add_argument("--num_data", type=int, required=True, help="Number of data points")
    p.add_argument("--num_queries", type=int, required=True, help="Number of queries")
    p.add_argument("--num_attrs", type=int, required=True, help="Number of attributes per point/query")
    p.add_argument("--min", type=float, required=True, help="Minimum attribute value")
    p.add_argument("--max", type=float, required=True, help="Maximum attribute value")
    p.add_argument("--minK", type=int, required=True, help="Minimum K for queries")
    p.add_argument("--maxK", type=int, required=True, help="Maximum K for queries")
    p.add_argument("--num_labels", type=int, required=True, help="Number of possible labels")
    p.add_argument("--output", type=str, required=True, help="Output file name")
    p.add_argument("--seed", type=int, default=42, help="Seed for RNG (default 42)")
    p.add_argument("--stream", choices=["python", "hash"], default="python",
                   help="python: reference-identical stream; hash: engine's counter-based stream")
    p.add_argument("--native", action="store_true", help="use tools/dmlp-gen for --stream hash")
    a = p.parse_args()

    if a.min >= a.max:
        sys.exit("Error: --min must be less than --max")
    if a.minK > a.maxK:
        sys.exit("Error: --minK must be <= --maxK")
    if a.num_labels <= 0:
        sys.exit("Error: --num_labels must be positive")

    if a.stream == "hash" and a.native:
        exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dmlp-gen")
        if os.path.exists(exe):
            subprocess.check_call([exe, str(a.num_data), str(a.num_queries), str(a.num_attrs), repr(a.min),
                                   repr(a.max), str(a.minK), str(a.maxK), str(a.num_labels), str(a.seed), a.output])
            print(f"Input file '{a.output}' generated (native).")
            return
    random.seed(a.seed)
    with open(a.output, "w") as f:
        (write_python_stream if a.stream == "python" else write_hash_stream)(f, a)
    print(f"Input file '{a.output}' generated successfully.")


if __name__ == "__main__":
    main()
