#!/usr/bin/env python
"""Small end-to-end case for `compute-sanitizer --tool memcheck|initcheck`: every kernel of the
library on sizes that exercise the ragged right edge, multi-strip rows, multi-domain pushes, the
flag/record protocol, early-exit replay, deep-halo epoch pairs and the text kernels."""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gol_b200  # noqa: E402
from gol_b200.ops import cpu, dense  # noqa: E402
from gol_b200.utils import patterns  # noqa: E402

import torch  # noqa: E402

for (w, h, grid) in [(2100, 96, (1, 1)), (1500, 120, (2, 2)), (840, 200, (1, 3))]:
    start = cpu.random_grid(w, h, 3)
    life = gol_b200.GameOfLife(w, h, grid=grid, device=0, gen_limit=40, batch_gens=16)
    life.load_packed(start)
    life.step(19)
    assert np.array_equal(life.packed_numpy(), cpu.step(start, w, 19))
    life.load_packed(start)
    res = life.run()
    want, wres = cpu.run(start, w, gen_limit=40)
    assert res.generations == wres.generations and np.array_equal(life.packed_numpy(), want)
    life.close()
# early exit + replay
cells = patterns.grid_with(256, 64, ("tromino", 100, 30))
life = gol_b200.GameOfLife(256, 64, device=0)
life.load(cells)
assert life.run().generations == 2
tmp = os.path.join(tempfile.gettempdir(), "san.txt")
life.save(tmp)
life.load_file(tmp)
life.close()
# deep-halo epoch pairs inside one process
os.environ["GOL_DEEP_HALO"] = "1"
for grid in [(1, 1), (2, 2)]:
    life = gol_b200.GameOfLife(1152, 300, device=0, grid=grid)
    start = cpu.random_grid(1152, 300, 5)
    life.load_packed(start)
    life.step(48)
    assert np.array_equal(life.packed_numpy(), cpu.step(start, 1152, 48))
    life.close()
del os.environ["GOL_DEEP_HALO"]
# dense ops
g = dense.random_grid(130, 33, 3, device="cuda:0")
f = torch.zeros(2, dtype=torch.int32, device="cuda:0")
o = dense.step(g, 130, flags=f)
t = dense.format_text(o, 130)
assert torch.equal(dense.parse_text(t, 130, 33), o)
assert dense.population(o, 130) == int(dense.unpack_cells(o, 130).sum().item())
print("sanitize case ok")
