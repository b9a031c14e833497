"""Model family: there is exactly one in the reference - B3/S23 on a torus - in six program
variants that differ only in parallelisation and I/O."""
from .life import GameOfLife, LifeConfig, VARIANT_OUTPUT  # noqa: F401


def __getattr__(name):
    if name == "LifeStep":   # imports torch lazily
        from .module import LifeStep
        return LifeStep
    raise AttributeError(name)
