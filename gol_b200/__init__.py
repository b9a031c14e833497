"""Importable alias of the ``game-of-life-in-parallel-mpi-openmp-cuda_b200`` package.

The product package directory carries the reference repository's name, which is not a valid
Python identifier; this shim makes it importable as ``gol_b200`` (sub-packages included:
``gol_b200.ops``, ``gol_b200.models``, ``gol_b200.parallel``, ``gol_b200.utils``).
"""
import os as _os

_PKG_DIR = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                         "game-of-life-in-parallel-mpi-openmp-cuda_b200")
__path__ = [_PKG_DIR]
with open(_os.path.join(_PKG_DIR, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_PKG_DIR, "__init__.py"), "exec"))
del _f
