"""Pattern fixtures: small named patterns stamped onto an empty grid.

Expected "Generations:" values on an otherwise empty torus (verified against the unmodified
serial reference): empty 0, single cell 1, block 2, L-tromino 2, blinker GEN_LIMIT, pulsar
GEN_LIMIT (period 3, consecutive generations always differ), glider GEN_LIMIT.
"""
from __future__ import annotations

import numpy as np

PATTERNS = {
    "cell": ["1"],
    "block": ["11", "11"],
    "tromino": ["11", "10"],            # becomes a block after one generation
    "blinker": ["111"],
    "toad": ["0111", "1110"],
    "beehive": ["0110", "1001", "0110"],  # still life
    "glider": ["010", "001", "111"],
    "rpentomino": ["011", "110", "010"],
    "pulsar": [
        "0011100011100", "0000000000000", "1000010100001", "1000010100001", "1000010100001",
        "0011100011100", "0000000000000", "0011100011100", "1000010100001", "1000010100001",
        "1000010100001", "0000000000000", "0011100011100",
    ],
    # a diagonal of three cells: dies in exactly two generations
    "diag3": ["100", "010", "001"],
    # "die hard": vanishes completely after 130 generations (empty exit deep inside a batch)
    "diehard": ["00000010", "11000000", "01000111"],
    # a 12 x 12 soup (found by search with the CPU oracle) that settles into still lifes only: on an empty
    # torus of at least 384 x 384 the reference loop reports Generations: 590 through the similarity exit
    "late590": ["110011011100", "001100001100", "100101000010", "111001111011", "010011100010", "001011001001",
                "110101011010", "110010101110", "001001011000", "110010101111", "100000100111", "110101010000"],
}


def pattern(name: str) -> np.ndarray:
    rows = PATTERNS[name]
    return np.array([[c == "1" for c in r] for r in rows], dtype=np.uint8)


def empty_grid(width: int, height: int) -> np.ndarray:
    return np.zeros((height, width), dtype=np.uint8)


def stamp(grid: np.ndarray, name_or_cells, x: int, y: int) -> np.ndarray:
    """Stamp a pattern with its top-left corner at (x, y); coordinates wrap around the torus."""
    p = pattern(name_or_cells) if isinstance(name_or_cells, str) else np.asarray(name_or_cells, dtype=np.uint8)
    h, w = grid.shape
    for dy in range(p.shape[0]):
        for dx in range(p.shape[1]):
            if p[dy, dx]:
                grid[(y + dy) % h, (x + dx) % w] = 1
    return grid


def grid_with(width: int, height: int, *stamps) -> np.ndarray:
    """grid_with(W, H, ("glider", x, y), ("block", x, y), ...)"""
    g = empty_grid(width, height)
    for name, x, y in stamps:
        stamp(g, name, x, y)
    return g


def conv_reference_step(cells: np.ndarray) -> np.ndarray:
    """Plain numpy torus step (roll-based), independent of all native code."""
    c = cells.astype(np.int32)
    n = sum(np.roll(np.roll(c, dy, 0), dx, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)) - c
    return ((n == 3) | ((n == 2) & (c == 1))).astype(np.uint8)
