"""gol_b200: Blackwell-native Conway's Game of Life engine.

Package layout
  models/    GameOfLife (engine-backed model), LifeConfig, reference program variants
  ops/       cpu (oracle, host text I/O) and dense (CUDA kernels on packed PyTorch tensors)
  parallel/  periodic R x C topology, decomposition, torch.distributed bootstrap, NCCL baseline
  utils/     pattern fixtures, reference builder, clock sampling, timers

The native library (libgolb200.so, C ABI in include/gol_b200.h) is built in-tree on first use.
"""
from . import _native  # noqa: F401
from ._native import GolError, build, cuda_available  # noqa: F401
from .ops import cpu  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    if name in ("GameOfLife", "LifeConfig", "VARIANT_OUTPUT"):
        from . import models
        return getattr(models, name)
    if name in ("models", "ops", "parallel", "utils"):
        import importlib
        return importlib.import_module(__name__ + "." + name)
    raise AttributeError(name)
