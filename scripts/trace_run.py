#!/usr/bin/env python
"""Per-launch device timeline of the generation loop (GOL_TRACE): kernel durations, gaps between
consecutive launches and flag spin time, for 1 GPU or under torch.distributed.run.

    python scripts/trace_run.py [--size 16384] [--grid RxC] [--out gpurun_out/trace]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("GOL_TRACE", "600")
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import gol_b200  # noqa: E402
from gol_b200.parallel import auto_grid  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=16384)
ap.add_argument("--width", type=int, default=0)
ap.add_argument("--height", type=int, default=0)
ap.add_argument("--grid", default="")
ap.add_argument("--out", default="gpurun_out/trace")
ap.add_argument("--step", type=int, default=0, help="time step(n) instead of run()")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
R, C = (int(v) for v in a.grid.split("x")) if a.grid else auto_grid(world)
W, H = (a.width or C * a.size), (a.height or R * a.size)
# warm-up engine run (module load, autotune) without a trace, then the traced one
for traced in (False, True):
    if not traced:
        saved = os.environ.pop("GOL_TRACE")
    else:
        os.environ["GOL_TRACE"] = saved
    life = gol_b200.GameOfLife(W, H, grid=(R, C), device=local)
    life.randomize(1)
    if a.step:
        ms = life.step(a.step)
    else:
        ms = life.run().loop_ms
    if traced:
        path = f"{a.out}_{R}x{C}_n{world}.rank{rank}.txt"
        life.trace_dump(path)
        rows = [list(map(int, l.split())) for l in open(path) if not l.startswith("#")]
        clk = 1.965  # cycles per ns at the max SM clock
        by_mode = {}
        for r in rows:
            by_mode.setdefault((r[1], r[2], r[3]), []).append(r)
        print(f"[rank {rank}] {W}x{H} on {R}x{C}: loop {ms:.3f} ms, {len(rows)} strip launches", flush=True)
        for (k, mode, fl), rs in sorted(by_mode.items()):
            dur = sum(r[5] for r in rs) / len(rs) / 1e3
            gap = sum(r[6] for r in rs) / len(rs) / 1e3
            spin = sum(r[8] for r in rs) / len(rs) / clk / 1e3
            print(f"[rank {rank}]   k={k} mode={mode} flags={fl}: n={len(rs)} mean duration {dur:.2f} us, mean gap before {gap:.2f} us, "
                  f"mean longest flag spin {spin:.2f} us", flush=True)
    life.close()
if world > 1:
    dist.destroy_process_group()
