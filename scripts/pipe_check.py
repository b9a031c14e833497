#!/usr/bin/env python
"""Correctness + timing of the warp-pipelined kernel (GOL_PIPE=LVxD must be set by the caller)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gol_b200  # noqa: E402
from gol_b200.ops import cpu  # noqa: E402

bad = 0
for (w, h, grid) in [(2100, 300, (1, 1)), (1500, 420, (2, 2)), (840, 200, (1, 1)), (4096, 1000, (1, 1))]:
    start = cpu.random_grid(w, h, 3)
    life = gol_b200.GameOfLife(w, h, grid=grid, device=0, gen_limit=100)
    life.load_packed(start)
    life.step(53)
    ok1 = np.array_equal(life.packed_numpy(), cpu.step(start, w, 53))
    life.load_packed(start)
    res = life.run()
    want, wres = cpu.run(start, w, gen_limit=100)
    ok2 = res.generations == wres.generations and np.array_equal(life.packed_numpy(), want)
    print(os.environ.get("GOL_PIPE"), w, h, grid, "step", ok1, "run", ok2, flush=True)
    bad += (not ok1) + (not ok2)
    life.close()
if bad:
    sys.exit(1)
print("pipe check ok", flush=True)
if "--no-timing" in sys.argv:
    sys.exit(0)
life = gol_b200.GameOfLife(16384, 16384, device=0)
life.randomize(1)
for i in range(2):
    ms = life.step(960)
print(f"{os.environ.get('GOL_PIPE')}: 16384^2 {ms * 1e3 / 960:.2f} us/generation, {16384 * 16384 * 960 / ms / 1e9:.2f} T/s", flush=True)
