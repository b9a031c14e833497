#!/usr/bin/env python
"""Cost of an early exit under the speculative driver (16384 x 16384, 1 GPU): the batch that overshoots,
the replay of the exiting batch one generation at a time, against the same number of generations with
no exit.  Prints one line per case for BASELINE.md."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gol_b200  # noqa: E402
from gol_b200.ops import cpu  # noqa: E402
from gol_b200.utils import patterns  # noqa: E402

n = 16384
life = gol_b200.GameOfLife(n, n, device=0)
for name in ("late590", "diehard"):
    start = cpu.pack(patterns.grid_with(n, n, (name, n // 2, n // 2)))
    for rep in range(2):
        life.load_packed(start)
        res = life.run()
    print(f"[early-exit] {name}: Generations {res.generations} ({res.exit_reason}), computed {res.generations_computed} generations "
          f"(speculation + replay), loop {res.loop_ms:.3f} ms = {res.loop_ms * 1e3 / max(1, res.generations):.2f} us per reported generation",
          flush=True)
life.randomize(3)
for g in (590, 130):
    ms = life.step(g)
    print(f"[early-exit] no exit: step({g}) {ms:.3f} ms = {ms * 1e3 / g:.2f} us per generation", flush=True)
life.close()
