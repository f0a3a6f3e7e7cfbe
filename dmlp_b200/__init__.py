"""Importable alias for the ``distributed-machine-learning-project_b200`` package.

The package directory carries the project's hyphenated name, which Python
cannot import directly; this shim points ``dmlp_b200`` at that directory so
``import dmlp_b200`` and ``import dmlp_b200.ops`` resolve to the real modules.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "distributed-machine-learning-project_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _os, _f
