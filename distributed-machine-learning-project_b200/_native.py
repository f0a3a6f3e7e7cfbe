"""Locate, build and load the in-tree native libraries.

``libdmlp_cpu.so``  CPU oracle, multi-"rank" strategies, parser, host generator
``libdmlp_knn.so``  sm_100a kernels and runtime (C ABI in include/dmlp_knn.h)

Both live in ``<package>/lib`` so they travel with the source tree and are
visibly loaded from it.  ``build()`` drives ``make libs``.
"""
import ctypes
import os
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
LIB_DIR = os.path.join(PKG_DIR, "lib")

_cache = {}


def build(targets=("libs",), quiet=True):
    """Run ``make`` for the given targets at the repo root."""
    cmd = ["make", "-C", ROOT, "-j8"] + list(targets)
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("native build failed:\n" + res.stdout[-4000:])
    if not quiet:
        print(res.stdout)
    return True


def _load(name, build_if_missing=True):
    if name in _cache:
        return _cache[name]
    path = os.path.join(LIB_DIR, name)
    if not os.path.exists(path) and build_if_missing:
        build((os.path.relpath(path, ROOT),))
    lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
    _cache[name] = lib
    return lib


def cpu_lib():
    return _load("libdmlp_cpu.so")


def knn_lib():
    """The CUDA library.  Raises if it is missing: GPU ops never fall back silently."""
    return _load("libdmlp_knn.so")


def torch_ops():
    """Load the ATen operator library (torch.ops.dmlp_b200.*) and return that namespace."""
    import torch
    if "torch_ops" not in _cache:
        knn_lib()  # its dependency, from the same directory
        path = os.path.join(LIB_DIR, "libdmlp_torch.so")
        if not os.path.exists(path):
            build((os.path.relpath(path, ROOT),))
        torch.ops.load_library(path)
        _cache["torch_ops"] = torch.ops.dmlp_b200
    return _cache["torch_ops"]
