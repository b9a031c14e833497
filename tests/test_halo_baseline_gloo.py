"""Multi-process CPU tests (gloo, world_size 2 and 4): the baseline halo-exchange protocol
(row phase, column phase over the padded height, corner transport, self-neighbour wrap, message
ordering when both neighbours are the same rank) against the global CPU oracle."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world, port, *extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "scripts", "halo_baseline_check.py"), "--backend", "gloo", *extra]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "halo baseline ok" in p.stdout


@pytest.mark.parametrize("world,grid,k", [(2, "1x2", 4), (2, "2x1", 3), (4, "2x2", 5), (4, "1x4", 2), (4, "4x1", 1)])
def test_halo_exchange_protocol(world, grid, k):
    _run(world, 29540 + world + k, "--width", "512", "--height", "96", "--k", str(k), "--epochs", "4", "--grid", grid)
