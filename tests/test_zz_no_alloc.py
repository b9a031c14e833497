"""GPU test of gol_engine_population_static, the allocation-free twin of gol_engine_population.

Who needs it: hosts of this library that must not call cudaMalloc themselves or through us - a
process whose GPU memory belongs to a framework pool (a raw cudaMalloc then fails or competes
with the pool), or a latency-bound loop that cannot take the device-wide stall of
cudaMalloc/cudaFree.  The entry point keeps its staging copy and its counter in a static
device array inside the library, so it is called here straight through the C ABI (ctypes),
twice in a row per grid: the second call only agrees with the first if the counter left by
the first one was cleared.  The file name sorts last on purpose, so a run that stops at the
first failure has already been through every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _entry(lib):
    fn = lib.gol_engine_population_static
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    return fn


def _both(lib, life):
    """(existing entry point, new entry point called twice) for the engine's current grid."""
    want = C.c_uint64()
    assert lib.gol_engine_population(life._h, C.byref(want)) == 0, lib.gol_last_error()
    got = []
    for _ in range(2):
        c = C.c_uint64(0xDEADBEEF)
        rc = _entry(lib)(life._h, C.byref(c))
        assert rc == 0, lib.gol_last_error().decode()
        got.append(c.value)
    return want.value, got


@pytest.mark.parametrize("w,h,grid", [(1100, 200, (1, 1)), (2240, 384, (2, 2)), (30, 30, (1, 1))])
def test_static_population_matches_allocating_entry_point(cuda_device, lib, w, h, grid):
    import gol_b200
    from gol_b200.ops import cpu
    start = cpu.random_grid(w, h, 17)
    life = gol_b200.GameOfLife(w, h, grid=grid, device=0)
    life.load_packed(start)
    want, got = _both(lib, life)
    assert want == int(cpu.unpack(start, w).sum())
    assert got == [want, want]          # exact: these are integer counts, the tests' tolerance is 0
    life.step(5)
    want2, got2 = _both(lib, life)
    assert want2 == int(cpu.unpack(cpu.step(start, w, 5), w).sum())
    assert got2 == [want2, want2]
    life.close()


def test_static_population_rejects_domains_beyond_the_static_maximum(cuda_device, lib):
    import gol_b200
    life = gol_b200.GameOfLife(8192, 8192, device=0)   # 8 MiB of cells > the 4 MiB staging area
    life.randomize(3)
    c = C.c_uint64()
    assert _entry(lib)(life._h, C.byref(c)) != 0
    msg = lib.gol_last_error().decode()
    assert "static scratch" in msg and "gol_engine_population" in msg
    # split into four domains each quarter fits, and the count agrees with the allocating path
    life.close()
    life = gol_b200.GameOfLife(8192, 8192, grid=(2, 2), device=0)
    life.randomize(3)
    want, got = _both(lib, life)
    assert got == [want, want] and abs(want / (8192 * 8192) - 0.5) < 0.01
    life.close()
