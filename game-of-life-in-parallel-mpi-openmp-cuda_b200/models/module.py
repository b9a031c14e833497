"""`LifeStep`: the B3/S23 torus step as a `torch.nn.Module` over uint8 cell tensors, for code that
wants to treat the automaton as a layer (batch of grids in, batch of grids out).  Each grid goes
through the packed CUDA kernels (pack -> n generations -> unpack); grids of at least 128 x 16 cells
use the strip engine, smaller ones the generic kernel.  No parameters, no autograd: Life is not
differentiable."""
from __future__ import annotations

import torch

from ..ops import dense


class LifeStep(torch.nn.Module):
    def __init__(self, generations: int = 1):
        super().__init__()
        self.generations = int(generations)

    @torch.no_grad()
    def forward(self, cells: torch.Tensor) -> torch.Tensor:
        """cells: uint8/bool CUDA tensor [..., H, W] (non-zero = alive) -> same shape, uint8."""
        if not cells.is_cuda:
            raise ValueError("LifeStep runs on CUDA tensors (use gol_b200.ops.cpu for the CPU oracle)")
        shape = cells.shape
        h, w = int(shape[-2]), int(shape[-1])
        flat = cells.reshape(-1, h, w).to(torch.uint8)
        out = torch.empty_like(flat)
        for i in range(flat.shape[0]):
            a = dense.pack_cells(flat[i].contiguous())
            b = torch.empty_like(a)
            for _ in range(self.generations):
                dense.step(a, w, out=b)
                a, b = b, a
            out[i] = dense.unpack_cells(a, w)
        return out.reshape(shape)
