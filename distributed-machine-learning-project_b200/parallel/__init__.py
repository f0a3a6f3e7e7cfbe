"""Parallel layer: process grid, distributed engine (fused and NCCL+cuBLAS baseline)."""
from .engine import DistributedKNN  # noqa: F401
from .grid import Grid, even_slice, reference_block  # noqa: F401
