"""Multi-process GPU tests: one process per GPU (CUDA IPC, fused peer pushes, release/acquire flags) for
every decomposition the box has GPUs for, and the failure path of the flag waits.  Ranks are never run as
several processes on ONE GPU: kernels of different processes that spin on each other are not guaranteed to
be co-scheduled there (the single-GPU stand-in is all domains in one launch, tests/test_gpu_engine.py)."""
import os
import random
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def torchrun(nproc, script, *args, timeout=600, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + random.randrange(300)), os.path.join(ROOT, "scripts", script), *args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)


def test_wait_timeout_is_reported_not_waited_for(cuda_device, monkeypatch):
    # fault injection (R16): GOL_FAULT_INJECT=wait makes the border warps of every epoch wait for a flag
    # sequence that no neighbour will ever publish - what a dead neighbouring rank looks like from
    # this rank.  The wait must give up after GOL_WAIT_TIMEOUT_MS, the engine must report it through
    # gol_last_error, nothing may hang, and the abandoned items must not be passed off as a result.
    # (One launch that spins on its own, no second kernel or process involved.)
    import time

    import gol_b200
    from gol_b200.ops import cpu
    monkeypatch.setenv("GOL_FAULT_INJECT", "wait")
    monkeypatch.setenv("GOL_WAIT_TIMEOUT_MS", "300")
    life = gol_b200.GameOfLife(1024, 512, grid=(2, 1), device=0)
    life.load_packed(cpu.random_grid(1024, 512, 2))
    t0 = time.time()
    with pytest.raises(gol_b200.GolError, match="wait timed out"):
        life.step(16)
    assert time.time() - t0 < 10
    life.close()
    monkeypatch.delenv("GOL_FAULT_INJECT")
    life = gol_b200.GameOfLife(1024, 512, grid=(2, 1), device=0)      # and a fresh engine is fine afterwards
    life.load_packed(cpu.random_grid(1024, 512, 2))
    life.step(16)
    life.close()


@pytest.mark.parametrize("grid", ["2x1", "1x2", "2x2", "4x1", "8x1", "1x8", "2x4"])
def test_one_process_per_gpu(cuda_device, grid):
    r, c = (int(v) for v in grid.split("x"))
    if torch.cuda.device_count() < r * c:
        pytest.skip(f"needs {r * c} GPUs")
    p = torchrun(r * c, "mgpu_check.py", "--grid", grid, "--size", "2048", "--gens", "32")
    out = p.stdout
    assert p.returncode == 0, out[-3000:] + p.stderr[-3000:]
    assert "steps bit-identical to the CPU oracle" in out and "byte-identical" in out and "early-exit semantics identical" in out
    assert "[check] checksum equals the single-process run" in out
