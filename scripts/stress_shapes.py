#!/usr/bin/env python
"""Extreme shapes and decompositions against the CPU oracle (one-off robustness sweep)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gol_b200  # noqa: E402
from gol_b200.ops import cpu  # noqa: E402

cases = [
    (128, 16, (1, 1)), (128, 200000, (1, 1)), (1000000, 16, (1, 1)), (100037, 130, (1, 1)), (70001, 257, (1, 1)),
    (8192, 2048, (4, 4)), (8192, 2048, (1, 16)), (4096, 4096, (16, 1)), (8256, 1000, (3, 5)), (1921, 64, (1, 1)),
    (3841, 64, (1, 1)), (1983, 999, (2, 3)), (129, 17, (1, 1)), (191, 16, (1, 1)), (255, 33, (1, 1)),
]
bad = 0
for w, h, grid in cases:
    t0 = time.time()
    start = cpu.random_grid(w, h, 9)
    try:
        life = gol_b200.GameOfLife(w, h, grid=grid, device=0, gen_limit=30)
    except Exception as e:  # noqa: BLE001
        print(f"{w}x{h} {grid}: engine refused: {str(e)[:90]}")
        continue
    life.load_packed(start)
    life.step(21)
    ok1 = np.array_equal(life.packed_numpy(), cpu.step(start, w, 21))
    life.load_packed(start)
    res = life.run()
    want, wres = cpu.run(start, w, gen_limit=30)
    ok2 = res.generations == wres.generations and np.array_equal(life.packed_numpy(), want) and res.population == wres.population
    life.close()
    print(f"{w}x{h} {grid}: step {'ok' if ok1 else 'MISMATCH'}, run {'ok' if ok2 else 'MISMATCH'} ({time.time() - t0:.1f}s)", flush=True)
    bad += (not ok1) + (not ok2)
print("stress shapes:", "all ok" if not bad else f"{bad} FAILURES")
sys.exit(1 if bad else 0)
