"""CPU ops: the threaded bit-packed oracle and host text conversion (numpy in / numpy out).

These wrap the native CPU implementation of the serial reference semantics
(``src/game.c:60-131, 169-196`` of the reference); they are the oracle for every device test
and the only part of the package that runs without a GPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from .. import _native as N


@dataclass
class RunResult:
    generations: int          # the value the reference prints as "Generations:"
    exit_reason: str          # "limit" | "empty" | "similar"
    generations_computed: int
    loop_ms: float
    population: int

    @classmethod
    def from_native(cls, r: N.Result) -> "RunResult":
        return cls(int(r.generations), N.EXIT_NAMES[int(r.exit_reason)], int(r.generations_computed),
                   float(r.loop_ms), int(r.population))


def words_per_row(width: int) -> int:
    return (width + 63) // 64


def pack(cells: np.ndarray) -> np.ndarray:
    """uint8/bool [H, W] (non-zero = alive) -> uint64 [H, ceil(W/64)], bit i of word q = cell 64q+i."""
    cells = np.asarray(cells)
    h, w = cells.shape
    nq = words_per_row(w)
    bits = np.zeros((h, nq * 64), dtype=np.uint8)
    bits[:, :w] = cells != 0
    return np.packbits(bits.reshape(h, nq, 64), axis=-1, bitorder="little").view(np.uint64).reshape(h, nq)


def unpack(packed: np.ndarray, width: int) -> np.ndarray:
    packed = np.ascontiguousarray(packed, dtype=np.uint64)
    h, nq = packed.shape
    bits = np.unpackbits(packed.view(np.uint8).reshape(h, nq * 8), axis=-1, bitorder="little")
    return bits[:, :width].copy()


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def step(packed: np.ndarray, width: int, generations: int = 1, threads: int = 0) -> np.ndarray:
    """Advance a packed grid by `generations` torus generations (no early exit)."""
    out = np.ascontiguousarray(packed, dtype=np.uint64).copy()
    N.check(N.lib().gol_cpu_step(_ptr(out), width, out.shape[0], generations, threads), "gol_cpu_step")
    return out


def run(packed: np.ndarray, width: int, gen_limit: int = 1000, similarity_frequency: int = 3,
        threads: int = 0):
    """The serial generation loop with both early exits; returns (final grid, RunResult)."""
    out = np.ascontiguousarray(packed, dtype=np.uint64).copy()
    res = N.Result()
    N.check(N.lib().gol_cpu_run(_ptr(out), width, out.shape[0], gen_limit, similarity_frequency, threads,
                                C.byref(res)), "gol_cpu_run")
    return out, RunResult.from_native(res)


def parse_text(text: bytes, width: int, height: int, strict: bool = True) -> np.ndarray:
    out = np.zeros((height, words_per_row(width)), dtype=np.uint64)
    N.check(N.lib().gol_text_parse(text, len(text), width, height, int(strict), _ptr(out)), "gol_text_parse")
    return out


def format_text(packed: np.ndarray, width: int) -> bytes:
    packed = np.ascontiguousarray(packed, dtype=np.uint64)
    h = packed.shape[0]
    buf = C.create_string_buffer(h * (width + 1))
    N.check(N.lib().gol_text_format(_ptr(packed), width, h, buf), "gol_text_format")
    return buf.raw


def read_file(path: str, width: int, height: int, strict: bool = True) -> np.ndarray:
    out = np.zeros((height, words_per_row(width)), dtype=np.uint64)
    N.check(N.lib().gol_file_read_packed(path.encode(), width, height, int(strict), _ptr(out)), "read")
    return out


def write_file(path: str, packed: np.ndarray, width: int) -> None:
    packed = np.ascontiguousarray(packed, dtype=np.uint64)
    N.check(N.lib().gol_file_write_packed(path.encode(), _ptr(packed), width, packed.shape[0]), "write")


def random_grid(width: int, height: int, seed: int = 1) -> np.ndarray:
    """Seeded i.i.d. P(1)=1/2 grid; identical to the device generator for the same seed."""
    out = np.zeros((height, words_per_row(width)), dtype=np.uint64)
    N.check(N.lib().gol_random_packed(_ptr(out), width, height, seed), "gol_random_packed")
    return out


def random_text_file(path: str, width: int, height: int, seed: int = 1) -> None:
    N.check(N.lib().gol_random_text_file(path.encode(), width, height, seed), "gol_random_text_file")


def write_pbm(path: str, packed: np.ndarray, width: int) -> None:
    packed = np.ascontiguousarray(packed, dtype=np.uint64)
    N.check(N.lib().gol_write_pbm(path.encode(), _ptr(packed), width, packed.shape[0]), "gol_write_pbm")


def grid_hash(packed: np.ndarray, width: int) -> int:
    """Position-dependent 64-bit checksum of a packed grid (gol_hash_packed); equals the sum of
    GameOfLife.local_hash() over all ranks for the same grid."""
    packed = np.ascontiguousarray(packed, dtype=np.uint64)
    c = C.c_uint64()
    N.check(N.lib().gol_hash_packed(packed.ctypes.data_as(C.c_void_p), width, packed.shape[0], C.byref(c)), "hash_packed")
    return int(c.value)
