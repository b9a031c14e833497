#!/usr/bin/env python
"""Timing control: N ranks under torchrun/NCCL, each running an INDEPENDENT single-domain engine
(no IPC mapping, no halo traffic) - separates process/NCCL effects from the fused exchange."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gol_b200  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
life = gol_b200.GameOfLife(16384, 16384, grid=(1, 1), device=local)
life.randomize(1)
for rep in range(3):
    dist.barrier()
    ms = life.step(256)
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"[indep] step(256) per rank 16384^2: {t.item():.3f} ms ({t.item() * 1e3 / 256:.2f} us/gen)", flush=True)
dist.destroy_process_group()
