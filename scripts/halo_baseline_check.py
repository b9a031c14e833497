#!/usr/bin/env python
"""Protocol check of the BASELINE halo exchange (torch.distributed P2P + all_reduce), one block per
rank.  Runs on CPUs with gloo (tests) or on GPUs with NCCL (timing against the fused path):

    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29533 \
        scripts/halo_baseline_check.py --backend gloo --width 512 --height 96 --k 4 --epochs 5
    ... --backend nccl --width 32768 --height 16384 --k 8 --epochs 32 --time
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gol_b200.ops import cpu  # noqa: E402
from gol_b200.parallel import Topology, auto_grid  # noqa: E402
from gol_b200.parallel.halo_baseline import EngineBlockGPU, PackedBlockCPU  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--backend", default="gloo")
ap.add_argument("--width", type=int, default=512)
ap.add_argument("--height", type=int, default=96)
ap.add_argument("--k", type=int, default=4)
ap.add_argument("--epochs", type=int, default=5)
ap.add_argument("--grid", default="")
ap.add_argument("--time", action="store_true")
a = ap.parse_args()

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", "0"))
if a.backend == "nccl":
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
else:
    dist.init_process_group("gloo")
R, C = (int(v) for v in a.grid.split("x")) if a.grid else auto_grid(world)
topo = Topology(a.width, a.height, R, C)

if a.backend == "nccl":
    blk = EngineBlockGPU(topo, rank, a.k, local)
else:
    blk = PackedBlockCPU(topo, rank, a.k)

if a.time:
    blk.randomize(1)
    for rep in range(3):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.epochs):
            blk.epoch()
        blk.flags_all()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gens = a.epochs * a.k
        if rank == 0:
            print(f"[baseline nccl] {a.width}x{a.height} on {R}x{C}, k={a.k}: {gens} generations in {t.item():.3f} ms -> "
                  f"{a.width * a.height * gens / t.item() / 1e9:.2f} T cell-updates/s ({t.item() * 1e3 / gens:.2f} us/gen), "
                  f"{blk.ex.messages} messages / {blk.ex.bytes} bytes sent by rank 0 so far", flush=True)
else:
    start = cpu.random_grid(a.width, a.height, 13)
    blk.load_global(start)
    want = start
    for e in range(a.epochs):
        blk.epoch()
        want = cpu.step(want, a.width, a.k)
        x0, y0, w, h = topo.extent(rank)
        got = blk.interior()
        ok = np.array_equal(got, want[y0:y0 + h, x0 // 64:(x0 + w) // 64])
        f = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda" if a.backend == "nccl" else "cpu")
        dist.all_reduce(f, op=dist.ReduceOp.MIN)
        if not f.item():
            raise SystemExit(f"rank {rank}: halo baseline mismatch after epoch {e + 1}")
    assert blk.flags_all() == bool(want.any())
    if rank == 0:
        print(f"halo baseline ok: {a.width}x{a.height} on {R}x{C}, k={a.k}, {a.epochs} epochs, backend {a.backend}", flush=True)
dist.destroy_process_group()
