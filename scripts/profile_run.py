#!/usr/bin/env python
"""Profiling driver: one run() of the reference loop on a random grid with a FIXED chunk height
(GOL_CHUNK_ROWS, so that no autotune launches precede the loop and launch indices are deterministic):
scan_alive, gather, then per 128-generation batch 15 x life_strip_kernel<8,false>, 1 x <8,true>, gather.

    GOL_CHUNK_ROWS=85 python scripts/profile_run.py [--size 16384] [--gens 256]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("GOL_CHUNK_ROWS", "85")
ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=16384)
ap.add_argument("--gens", type=int, default=256)
ap.add_argument("--grid", default="1x1")
a = ap.parse_args()

import gol_b200  # noqa: E402

r, c = (int(v) for v in a.grid.split("x"))
life = gol_b200.GameOfLife(a.size, a.size, grid=(r, c), device=0, gen_limit=a.gens)
life.randomize(1)
res = life.run()
print(f"run: {res.generations} generations of {a.size}^2 in {res.loop_ms:.3f} ms, chunk_rows {life.chunk_rows}, "
      f"checksum {life.local_hash():016x}, population {life.population()}", flush=True)
life.close()
