"""Operators: ``cpu`` (oracle and host utilities) and ``cuda`` (sm_100a kernels)."""
from . import cpu  # noqa: F401
