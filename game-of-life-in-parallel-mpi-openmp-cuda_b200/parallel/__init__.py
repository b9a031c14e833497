"""Parallel layer: topology / decomposition, launch helpers and the NCCL halo baseline."""
from .topology import DIRECTIONS, OPPOSITE, Topology, auto_grid, near_square_grid  # noqa: F401
