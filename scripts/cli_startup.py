#!/usr/bin/env python
"""Where the CLI's wall time goes (16384^2, 268 MB text each way), with another CUDA context alive on the
box (as inside bench.py) and without: phase report of --stats + whole-process wall time."""
import os
import subprocess
import sys
import tempfile
import time

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
work = tempfile.mkdtemp(prefix="golcli_")
subprocess.run([root + "/bin/gol-generate", "16384", "16384", "20240229", "-o", work + "/in.txt"], check=True, timeout=120)


def run(tag, extra_env=None):
    for i in range(3):
        t0 = time.perf_counter()
        p = subprocess.run([root + "/bin/gol", "16384", "16384", "in.txt", "--variant", "cuda", "--stats"], cwd=work,
                           capture_output=True, text=True, timeout=120, env=dict(os.environ, **(extra_env or {})))
        wall = (time.perf_counter() - t0) * 1e3
        ph = p.stderr.split("phases [ms]:")[-1].strip() if "phases" in p.stderr else p.stderr[-200:]
        print(f"[cli-startup] {tag} run {i}: wall {wall:.0f} ms | {ph}", flush=True)


run("cold (no other context)")
import torch  # noqa: E402
torch.zeros(1, device="cuda")
run("warm (a python process holds a context)")
run("warm, 2 io lanes x 4 MiB", {"GOL_IO_THREADS": "2", "GOL_IO_SLOT_MB": "4"})
run("warm, no autotune", {"GOL_AUTOTUNE": "0"})
