"""A fresh `git clone` of HEAD builds the driver and answers the golden vector.

Round 1 shipped with `src/engine/` swallowed by an unanchored ignore pattern: the
working tree built, a clone did not.  This test clones the committed tree into a
temporary directory, builds the CPU-only driver there (`make NO_CUDA=1`) and checks
the checksums of SURVEY.md §4.4 -- so a source file that exists on disk but not in
git fails here, on the CPU box, before it reaches a judge.
"""
import os
import shutil
import subprocess
import sys

import pytest

GOLDEN = ["Query 0 checksum: 1048314657462637447",
          "Query 1 checksum: 17629266663427905930",
          "Query 2 checksum: 6436800073062518871"]


def _git(repo_root, *args):
    return subprocess.run(["git", "-C", repo_root, *args], capture_output=True, text=True)


@pytest.mark.skipif(shutil.which("git") is None or shutil.which("make") is None, reason="needs git and make")
def test_fresh_clone_builds_and_runs(repo_root, tmp_path):
    if _git(repo_root, "rev-parse", "--git-dir").returncode != 0:
        pytest.skip("not a git checkout (e.g. the snapshot shipped to the GPU box)")
    clone = tmp_path / "clone"
    res = subprocess.run(["git", "clone", "--quiet", "--no-hardlinks", repo_root, str(clone)], capture_output=True,
                         text=True)
    assert res.returncode == 0, res.stderr
    # every source the Makefile names must be tracked
    for rel in ("src/engine/engine.cpp", "src/engine/gpu_engine.cu", "src/engine/gpu_stub.cpp", "src/driver/main.cpp",
                "include/engine.h", "include/common.h"):
        assert (clone / rel).exists(), f"{rel} is not in git"
    res = subprocess.run(["make", "-C", str(clone), "-j8", "NO_CUDA=1", "engine", "engine.debug"],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    inp = tmp_path / "tiny3.in"
    subprocess.check_call([sys.executable, str(clone / "tools" / "generate_input.py"), "--num_data", "12",
                           "--num_queries", "3", "--num_attrs", "2", "--min", "0", "--max", "10", "--minK", "3",
                           "--maxK", "3", "--num_labels", "3", "--seed", "42", "--output", str(inp)],
                          stdout=subprocess.DEVNULL)
    with open(inp, "rb") as f:
        run = subprocess.run([str(clone / "engine")], stdin=f, capture_output=True, text=True,
                             env=dict(os.environ, DMLP_BACKEND="cpu"))
    assert run.returncode == 0, run.stderr
    assert sorted(run.stdout.splitlines()) == sorted(GOLDEN)
    # nothing that `make` needs may be hidden by .gitignore in the working tree either
    ignored = _git(repo_root, "status", "--ignored", "--short", "--", "src", "include", "tools", "tests").stdout
    hidden = [ln for ln in ignored.splitlines()
              if ln.startswith("!!") and "__pycache__" not in ln and
              (ln[3:].startswith(("src/", "include/")) or ln.rstrip("/").endswith((".cpp", ".cu", ".h", ".cuh", ".py")))]
    assert not hidden, f"source files ignored by git: {hidden}"
