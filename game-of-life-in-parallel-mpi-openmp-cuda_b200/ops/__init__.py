"""Ops: ``cpu`` (threaded bit-packed oracle, host text I/O) and ``dense`` (CUDA kernels on
packed PyTorch tensors).  ``dense`` imports torch lazily so the CPU ops stay usable without it."""
from . import cpu  # noqa: F401


def __getattr__(name):
    if name == "dense":
        import importlib
        return importlib.import_module(__name__ + ".dense")
    raise AttributeError(name)
