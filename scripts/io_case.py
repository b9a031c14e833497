#!/usr/bin/env python
"""Exercises the non-stencil kernels at a profiling-friendly size (8192 x 8192): text parse / format,
cell pack / unpack, generator, population, checksum, generic step, first halo exchange, alive scan + gather."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import gol_b200  # noqa: E402
from gol_b200.ops import dense  # noqa: E402

n = 8192
g = dense.random_grid(n, n, 3, device="cuda:0")
t = dense.format_text(g, n)
b = dense.parse_text(t, n, n)
assert torch.equal(b, g)
c = dense.unpack_cells(g, n)
assert torch.equal(dense.pack_cells(c), g)
f = torch.zeros(2, dtype=torch.int32, device="cuda:0")
o = dense.step(g, n, flags=f)
print("population", dense.population(o, n), "flags", f.tolist())
life = gol_b200.GameOfLife(n, n, device=0, gen_limit=16)
path = os.path.join(tempfile.gettempdir(), "io_case.txt")
life.load_packed(g)
life.save(path)
life.load_file(path)
print(life.run())
print("checksum", hex(life.local_hash()), "population", life.population())
life.close()
