#!/usr/bin/env python
"""Short driver for profiling: advances a random grid by a few epochs of the strip kernel.

    python tools/profile_step.py [--size 16384] [--gens 64] [--k 4] [--reps 3]

Prints device-timed cell-updates/s per repetition (CUDA events inside the engine)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=16384)
ap.add_argument("--height", type=int, default=0)
ap.add_argument("--gens", type=int, default=64)
ap.add_argument("--k", type=int, default=0)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--grid", default="1x1")
a = ap.parse_args()

import gol_b200  # noqa: E402

h = a.height or a.size
r, c = (int(v) for v in a.grid.split("x"))
life = gol_b200.GameOfLife(a.size, h, k_max=a.k, grid=(r, c), device=0)
life.randomize(1)
for i in range(a.reps):
    ms = life.step(a.gens)
    print(f"rep {i}: {a.gens} generations of {a.size}x{h} k={a.k or 'auto'}: {ms:.3f} ms "
          f"-> {a.size * h * a.gens / ms / 1e9:.2f} T cell-updates/s, {ms * 1e3 / a.gens:.2f} us/generation")
life.close()
