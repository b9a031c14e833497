#!/usr/bin/env python
"""Multi-GPU validation + timing, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 scripts/mgpu_check.py [--size 16384] [--gens 256]

Checks the fused P2P halo push / flag protocol against the CPU oracle (steps, early exits,
parallel file I/O), then times weak-scaling steps (device time, max over ranks)."""
import argparse
import os
import sys
import tempfile

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gol_b200  # noqa: E402
from gol_b200.ops import cpu  # noqa: E402
from gol_b200.parallel import auto_grid  # noqa: E402
from gol_b200.utils import patterns  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=16384)
ap.add_argument("--gens", type=int, default=256)
ap.add_argument("--skip-check", action="store_true")
ap.add_argument("--grid", default="")
ap.add_argument("--width", type=int, default=0, help="global width override (strong scaling configs)")
ap.add_argument("--height", type=int, default=0)
a = ap.parse_args()

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
R, C = (int(v) for v in a.grid.split("x")) if a.grid else auto_grid(world)


def local_equal(life, want):
    got = life.packed_numpy()
    ok = True
    for g in life.local_domains:
        x0, y0, w, h = life.domain_extent(g)
        q0, q1 = x0 // 64, (x0 + w + 63) // 64
        ok &= bool(np.array_equal(got[y0:y0 + h, q0:q1], want[y0:y0 + h, q0:q1]))
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def log(*args):
    if rank == 0:
        print(*args, flush=True)


if not a.skip_check:
    w, h = 4480, 1536
    start = cpu.random_grid(w, h, 5)
    life = gol_b200.GameOfLife(w, h, grid=(R, C), device=local)
    life.load_packed(start)
    want = start
    for n in (1, 5, 16, 37):
        life.step(n)
        want = cpu.step(want, w, n)
        assert local_equal(life, want), f"mismatch after +{n}"
    log(f"[check] {w}x{h} on {R}x{C}: steps bit-identical to the CPU oracle")
    # parallel file I/O
    tmp = tempfile.gettempdir()
    inp, out = os.path.join(tmp, "mg_in.txt"), os.path.join(tmp, "mg_out.txt")
    if rank == 0:
        cpu.random_text_file(inp, w, h, 9)
    dist.barrier()
    life.load_file(inp)
    assert local_equal(life, cpu.random_grid(w, h, 9))
    life.step(3)
    life.save(out)
    if rank == 0:
        with open(out, "rb") as f:
            assert f.read() == cpu.format_text(cpu.step(cpu.random_grid(w, h, 9), w, 3), w)
    log("[check] parallel read + collective-format write byte-identical")
    life.close()
    # early exits
    w, h = 1024, 512
    for name, stamps in {"empty": [], "block": [("block", 511, 255)], "tromino": [("tromino", 510, 254)],
                         "blinker": [("blinker", 600, 300)], "glider": [("glider", 1020, 508)],
                         "mix": [("tromino", 100, 100), ("diag3", 700, 400)]}.items():
        cells = patterns.grid_with(w, h, *stamps)
        startp = cpu.pack(cells)
        for kw in ({"gen_limit": 200}, {"gen_limit": 100, "similarity_frequency": 5, "batch_gens": 32}):
            wantg, wres = cpu.run(startp, w, **{k: v for k, v in kw.items() if k != "batch_gens"})
            life = gol_b200.GameOfLife(w, h, grid=(R, C), device=local, **kw)
            life.load_packed(startp)
            res = life.run()
            assert (res.generations, res.exit_reason) == (wres.generations, wres.exit_reason), (name, res, wres)
            assert local_equal(life, wantg), name
            life.close()
    log("[check] early-exit semantics identical to the serial loop on every rank")

# weak-scaling timing: size x size cells per GPU
W, H = (a.width or C * a.size), (a.height or R * a.size)
life = gol_b200.GameOfLife(W, H, grid=(R, C), device=local, similarity_frequency=3, gen_limit=1000)
life.randomize(1)
for rep in range(3):
    ms = life.step(a.gens)
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    log(f"[time] step({a.gens}) {W}x{H} on {R}x{C}: {t.item():.3f} ms -> "
        f"{W * H * a.gens / t.item() / 1e9:.2f} T cell-updates/s ({t.item() * 1e3 / a.gens:.2f} us/gen)")
life.randomize(1)
res = life.run()
t = torch.tensor([res.loop_ms], device="cuda", dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
log(f"[time] run() {res.generations} generations: {t.item():.3f} ms -> "
    f"{W * H * res.generations_computed / t.item() / 1e9:.2f} T cell-updates/s "
    f"(grid {R}x{C}, chunk_rows {life.chunk_rows}, deep halo {life.uses_deep_halo})")
# identity of the final grid: checksum summed over the ranks == the same grid run inside one process
lh = life.local_hash()
ident = torch.tensor([lh - (1 << 64) if lh >= (1 << 63) else lh, life.population()], dtype=torch.int64, device="cuda")
dist.all_reduce(ident, op=dist.ReduceOp.SUM)
life.close()
if not a.skip_check and W * H <= (1 << 34):
    if rank == 0:
        one = gol_b200.GameOfLife(W, H, grid=(1, 1), device=local, similarity_frequency=3, gen_limit=1000)
        one.randomize(1)
        r1 = one.run()
        same = (one.local_hash() == (int(ident[0].item()) & ((1 << 64) - 1)) and one.population() == int(ident[1].item())
                and r1.generations == res.generations)
        one.close()
        assert same, "multi-process result differs from the single-process run"
        log("[check] checksum equals the single-process run of the same grid")
    dist.barrier()
dist.destroy_process_group()
