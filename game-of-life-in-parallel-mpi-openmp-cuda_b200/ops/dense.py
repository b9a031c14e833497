"""Device ops on packed dense CUDA tensors (thin PyTorch surface over the gol_op_* kernels).

A packed grid is an ``int64`` CUDA tensor of shape ``[H, ceil(W/64)]`` (bit i of word q of a
row is cell ``64*q + i``; bits beyond the width are zero).  All ops run on the current
PyTorch CUDA stream and raise ``GolError`` if the native library has no usable device — there
is deliberately no eager fallback.

Reference counterparts: ``step`` = evolve + halo_rows + halo_cols (src/game_cuda.cu:52-74,
128-148), ``flags``/``empty``/``similar`` = empty + compare (:76-126), ``parse_text`` /
``format_text`` = the fgetc reader and fprintf writer (src/game.c:25-40, 149-167).
"""
from __future__ import annotations

import torch

from .. import _native as N


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _check_packed(t: torch.Tensor, width: int) -> None:
    if not t.is_cuda or t.dtype != torch.int64 or t.dim() != 2 or not t.is_contiguous():
        raise ValueError("packed grid must be a contiguous 2-D int64 CUDA tensor")
    if t.shape[1] != (width + 63) // 64:
        raise ValueError(f"packed grid has {t.shape[1]} words per row, width {width} needs {(width + 63) // 64}")


def words_per_row(width: int) -> int:
    return (width + 63) // 64


def step(packed: torch.Tensor, width: int, out: torch.Tensor | None = None,
         flags: torch.Tensor | None = None) -> torch.Tensor:
    """One torus generation of any size >= 1x1.  `flags` (int32[2], CUDA) is OR-ed with
    {any cell alive in out, out != packed}."""
    N.require_cuda()
    _check_packed(packed, width)
    if out is None:
        out = torch.empty_like(packed)
    with torch.cuda.device(packed.device):
        N.check(N.lib().gol_op_step_dense(packed.data_ptr(), out.data_ptr(), width, packed.shape[0],
                                          flags.data_ptr() if flags is not None else None, _stream()), "step")
    return out


def flags(a: torch.Tensor, width: int, b: torch.Tensor | None = None) -> tuple[bool, bool]:
    """(any cell alive in a, a differs from b) in one pass."""
    N.require_cuda()
    _check_packed(a, width)
    f = torch.zeros(2, dtype=torch.int32, device=a.device)
    with torch.cuda.device(a.device):
        N.check(N.lib().gol_op_flags_dense(a.data_ptr(), b.data_ptr() if b is not None else None, width,
                                           a.shape[0], f.data_ptr(), _stream()), "flags")
    alive, diff = f.tolist()
    return bool(alive), bool(diff)


def empty(a: torch.Tensor, width: int) -> bool:
    """True iff every cell is dead (reference `empty`)."""
    return not flags(a, width)[0]


def similar(a: torch.Tensor, b: torch.Tensor, width: int) -> bool:
    """True iff the two generations are identical (reference `similarity` / `compare`)."""
    return not flags(a, width, b)[1]


def pack_cells(cells: torch.Tensor) -> torch.Tensor:
    """uint8 [H, W] CUDA tensor (non-zero = alive) -> packed int64 [H, ceil(W/64)]."""
    N.require_cuda()
    if not cells.is_cuda or cells.dtype != torch.uint8 or cells.dim() != 2:
        raise ValueError("cells must be a 2-D uint8 CUDA tensor")
    cells = cells.contiguous()
    h, w = cells.shape
    out = torch.empty((h, words_per_row(w)), dtype=torch.int64, device=cells.device)
    with torch.cuda.device(cells.device):
        N.check(N.lib().gol_op_pack_cells(cells.data_ptr(), out.data_ptr(), w, h, _stream()), "pack_cells")
    return out


def unpack_cells(packed: torch.Tensor, width: int) -> torch.Tensor:
    N.require_cuda()
    _check_packed(packed, width)
    h = packed.shape[0]
    out = torch.empty((h, width), dtype=torch.uint8, device=packed.device)
    with torch.cuda.device(packed.device):
        N.check(N.lib().gol_op_unpack_cells(packed.data_ptr(), out.data_ptr(), width, h, _stream()), "unpack_cells")
    return out


def parse_text(text: torch.Tensor, width: int, height: int, pitch: int | None = None,
               strict: bool = True) -> torch.Tensor:
    """Fixed-record text (uint8 CUDA tensor, `pitch` bytes per line, default width+1) -> packed."""
    N.require_cuda()
    pitch = width + 1 if pitch is None else pitch
    if text.numel() < (height - 1) * pitch + width:
        raise ValueError("text buffer is too short")
    out = torch.empty((height, words_per_row(width)), dtype=torch.int64, device=text.device)
    err = torch.zeros(1, dtype=torch.int32, device=text.device)
    with torch.cuda.device(text.device):
        N.check(N.lib().gol_op_parse_text_n(text.data_ptr(), text.numel(), out.data_ptr(), width, height, pitch,
                                            err.data_ptr(), _stream()), "parse_text")
    code = int(err.item())
    if code & 2:
        raise N.GolError("parse_text: a line does not end in a newline at column `width`")
    if (code & 1) and strict:
        raise N.GolError("parse_text: invalid byte, only '0', '1' and newlines are accepted")
    return out


def format_text(packed: torch.Tensor, width: int, newline: bool = True) -> torch.Tensor:
    """packed -> uint8 CUDA tensor [H, width+1] of '0'/'1' + '\\n' (the reference file format)."""
    N.require_cuda()
    _check_packed(packed, width)
    h = packed.shape[0]
    pitch = width + (1 if newline else 0)
    out = torch.empty((h, pitch), dtype=torch.uint8, device=packed.device)
    with torch.cuda.device(packed.device):
        N.check(N.lib().gol_op_format_text(packed.data_ptr(), out.data_ptr(), width, h, pitch, int(newline),
                                           _stream()), "format_text")
    return out


def random_grid(width: int, height: int, seed: int = 1, device="cuda", x0: int = 0, y0: int = 0) -> torch.Tensor:
    """Seeded i.i.d. P(1)=1/2 grid (generate.sh replacement); bit-identical to ops.cpu.random_grid."""
    N.require_cuda()
    out = torch.empty((height, words_per_row(width)), dtype=torch.int64, device=device)
    with torch.cuda.device(out.device):
        N.check(N.lib().gol_op_random_dense(out.data_ptr(), width, height, x0, y0, seed, _stream()), "random")
    return out


def population(packed: torch.Tensor, width: int) -> int:
    N.require_cuda()
    _check_packed(packed, width)
    c = torch.zeros(1, dtype=torch.int64, device=packed.device)
    with torch.cuda.device(packed.device):
        N.check(N.lib().gol_op_popcount_dense(packed.data_ptr(), width, packed.shape[0], c.data_ptr(), _stream()),
                "population")
    return int(c.item())


# ---------------------------------------------------------------- torch.library registration
# The step is also registered as a PyTorch custom op (torch.ops.gol_b200.step_dense) with a fake
# (meta) implementation, so it composes with code that traces or inspects ops; the kernel behind
# it is the same hand-written CUDA kernel - there is no compiler on the hot path.
try:
    @torch.library.custom_op("gol_b200::step_dense", mutates_args=(), device_types="cuda")
    def step_dense_op(packed: torch.Tensor, width: int) -> torch.Tensor:
        return step(packed, width)

    @step_dense_op.register_fake
    def _(packed, width):
        return torch.empty_like(packed)
except Exception:  # pragma: no cover - very old torch without torch.library.custom_op
    step_dense_op = None
