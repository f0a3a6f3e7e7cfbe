"""Reference arm of bench.py: the UNMODIFIED reference (baseline/_ref/engine.cpp +
common.cpp) through its own public path -- text on stdin, Engine::KNN, `Time taken`
on stderr (reference: common.cpp:81-135).

The reference is a Makefile project that needs mpicxx; the image has no MPI, so
the sources are compiled against an MPI stand-in (baseline/mpi_shim/mpi.h): the handful
of calls the reference makes, for one rank or for several OS processes on one machine
joined by AF_UNIX sockets.  None of this repository's engine, kernels or models is on
this path: only the input generator (the text the reference parses) is ours, as it must be.

Several ranks (bench.py --impl reference under torch.distributed.run): every launched
process starts one reference rank (RANK / WORLD_SIZE from the launcher), rank 0 feeds
the text.  The reference only survives square process grids: on 2 or 8 ranks
(MPI_Dims_create -> 2x1, 4x2) it overruns its gather buffer (engine.cpp:275, SURVEY.md
§2.5 B1) and aborts with heap corruption -- observed here as "malloc(): invalid size" --
so those sizes report "unavailable"; 1, 4 and 9 ranks run and print the same results as
our engine.

The named config (num_attrs 8192, 65536 queries per GPU against 65536 points)
takes the reference about one second PER QUERY on one core, i.e. most of a day per
step, so the arm times a short run: the full dataset shape with `--ref-queries`
queries.  The reference scans the whole dataset for every query, so its queries/s
does not depend on how many queries are timed.  k is held at 16 because the
reference mis-indexes its lists when k varies.
"""
import json
import os
import re
import shutil
import subprocess
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.path.join(ROOT, "baseline", "_ref")
REF_BIN = os.path.join(ROOT, "baseline", "_ref_build", "engine_ref")


def ensure_reference():
    """Copy the reference sources (if reachable) and build them with the MPI stand-in."""
    if not os.path.exists(os.path.join(REF_SRC, "engine.cpp")):
        if os.path.exists("/root/reference/engine.cpp"):
            shutil.copytree("/root/reference", REF_SRC, dirs_exist_ok=True)
        else:
            return "reference sources not present under baseline/_ref"
    newest_src = max(os.path.getmtime(os.path.join(REF_SRC, f)) for f in ("engine.cpp", "common.cpp"))
    if not os.path.exists(REF_BIN) or os.path.getmtime(REF_BIN) < newest_src:
        os.makedirs(os.path.dirname(REF_BIN), exist_ok=True)
        tmp = REF_BIN + ".tmp%d" % os.getpid()
        cmd = ["g++", "-O3", "-I", os.path.join(ROOT, "baseline", "mpi_shim"), os.path.join(REF_SRC, "engine.cpp"),
               os.path.join(REF_SRC, "common.cpp"), "-o", tmp]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            return "reference does not compile here: " + res.stderr.strip().splitlines()[-1][:200]
        os.replace(tmp, REF_BIN)   # other launched ranks may be waiting for the binary: appear atomically
    return None


def _dims_create(n):
    """MPI_Dims_create(n, 2): balanced, non-increasing."""
    best = 1
    f = 1
    while f * f <= n:
        if n % f == 0:
            best = f
        f += 1
    return n // best, best


def _run_ranks(args, world, rank, out, err):
    """One reference rank per launched process (square grids only, see the module text)."""
    import tempfile
    port = os.environ.get("MASTER_PORT", "0")
    sock_dir = os.path.join(tempfile.gettempdir(), "dmlp_mpi_%s" % port)
    os.makedirs(sock_dir, exist_ok=True)
    gen = os.path.join(ROOT, "tools", "dmlp-gen")
    if rank == 0:
        if err is None and not os.path.exists(gen):
            if subprocess.run(["make", "-C", ROOT, "tools"], capture_output=True, text=True).returncode != 0:
                err = "input generator did not build"
        if err:
            # the other launched processes give up on their own: their rank finds no rank 0 to connect to
            out["unavailable"] = err
            print(json.dumps(out))
            return 0
    else:
        t0 = time.time()
        while not os.path.exists(REF_BIN):       # rank 0 may still be building it
            if time.time() - t0 > 300:
                return 0
            time.sleep(0.1)
    env = dict(os.environ, DMLP_MPI_DIR=sock_dir)
    n, a = args.num_data, args.num_attrs
    q = max(1, getattr(args, "ref_queries", 8)) * world      # weak scaling: the timed queries grow with the ranks
    if rank != 0:
        subprocess.run([REF_BIN], stdin=subprocess.DEVNULL, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, env=env)
        return 0
    gen_cmd = [gen, str(n), str(q), str(a), "-100", "100", "16", "16", str(args.labels), str(args.seed), "/dev/stdout"]
    t0 = time.time()
    gp = subprocess.Popen(gen_cmd, stdout=subprocess.PIPE)
    rp = subprocess.Popen([REF_BIN], stdin=gp.stdout, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, env=env)
    gp.stdout.close()
    _, stderr = rp.communicate()
    gp.wait()
    wall = time.time() - t0
    m = re.search(r"Time taken:\s*([0-9]+)\s*ms", stderr or "")
    if rp.returncode != 0 or not m:
        out["unavailable"] = "reference run failed: " + (stderr or "").strip()[-200:]
        print(json.dumps(out))
        return 0
    ms = max(float(m.group(1)), 1e-3)
    value = q / (ms / 1e3)
    rows, cols = _dims_create(world)
    out.update({
        "metric": "queries_per_sec (samples/sec of the whole k-NN job)", "value": value, "unit": "queries/s",
        "n_gpus": world, "steps": 1, "warmup": 0, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp64 (CPU, %d ranks)" % world,
        "data": "synthetic (generate_input.py distribution), text through the reference's own stdin parser",
        "config": {"model": "knn synthetic hidden(num_attrs)=%d" % a, "num_data_total": n, "seq_len": a,
                   "queries_timed": q, "k": [16, 16],
                   "note": "CPU ranks (the reference has no GPU code), one per launched process; its own %dx%d grid; "
                           "Time taken includes its scatter / broadcast of the inputs, as in its own driver" % (rows, cols),
                   "parallelism": "%d ranks (MPI stand-in over AF_UNIX sockets), grid %dx%d" % (world, rows, cols)},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "CPU only: the reference's timed region already is its end-to-end call"},
        "gpu_launches": 0, "wall_seconds_including_generation_and_parse": wall,
    })
    print(json.dumps(out))
    return 0


def run_reference(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    out = {"impl": "reference"}
    rows, cols = _dims_create(world)
    if rows != cols:
        if rank == 0:
            out["unavailable"] = ("the unmodified reference aborts on %d ranks: MPI_Dims_create gives a %dx%d grid and "
                                  "engine.cpp:275 sizes its gather buffer for a square one (heap corruption, "
                                  "'malloc(): invalid size'); it runs on 1, 4, 9, ... ranks" % (world, rows, cols))
            print(json.dumps(out))
        return 0
    err = ensure_reference() if rank == 0 else None
    if world > 1:
        return _run_ranks(args, world, rank, out, err)
    if rank != 0:
        return 0
    gen = os.path.join(ROOT, "tools", "dmlp-gen")
    if err is None and not os.path.exists(gen):
        res = subprocess.run(["make", "-C", ROOT, "tools"], capture_output=True, text=True)
        if res.returncode != 0:
            err = "input generator did not build"
    if err:
        out["unavailable"] = err
        print(json.dumps(out))
        return 0

    n, a = args.num_data, args.num_attrs
    q = max(1, getattr(args, "ref_queries", 8))
    steps, warmup = max(1, min(args.steps, 1)), 0
    # generator -> pipe -> reference stdin (a 5 GB text file never touches the disk)
    gen_cmd = [gen, str(n), str(q), str(a), "-100", "100", "16", "16", str(args.labels), str(args.seed), "/dev/stdout"]
    t0 = time.time()
    gp = subprocess.Popen(gen_cmd, stdout=subprocess.PIPE)
    rp = subprocess.Popen([REF_BIN], stdin=gp.stdout, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
    gp.stdout.close()
    _, stderr = rp.communicate()
    gp.wait()
    wall = time.time() - t0
    m = re.search(r"Time taken:\s*([0-9]+)\s*ms", stderr or "")
    if rp.returncode != 0 or not m:
        out["unavailable"] = "reference run failed: " + (stderr or "").strip()[-200:]
        print(json.dumps(out))
        return 0
    ms = max(float(m.group(1)), 1e-3)
    value = q / (ms / 1e3)
    out.update({
        "metric": "queries_per_sec (samples/sec of the whole k-NN job)", "value": value, "unit": "queries/s",
        "n_gpus": 1, "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp64 (CPU, one rank)",
        "data": "synthetic (generate_input.py distribution), text through the reference's own stdin parser",
        "config": {"model": "knn synthetic hidden(num_attrs)=%d" % a, "num_data_total": n, "seq_len": a,
                   "queries_timed": q, "k": [16, 16],
                   "note": "same dataset shape as the GPU arm; the reference needs ~1 s per query at this shape, "
                           "so %d queries are timed (its cost per query does not depend on the query count); "
                           "Time taken excludes the reference's text parse, as in its own driver" % q,
                   "parallelism": "1 rank (MPI stand-in)"},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "CPU only: the reference's timed region already is its end-to-end call"},
        "gpu_launches": 0, "wall_seconds_including_generation_and_parse": wall,
    })
    print(json.dumps(out))
    return 0
