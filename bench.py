#!/usr/bin/env python
"""Headline benchmark: cell-updates/sec of the reference generation loop.

Config (BASELINE.json #2): 16384 x 16384 random torus per GPU, GEN_LIMIT = 1000 generations,
similarity check every 3 generations (the reference defaults).  One *step* is one complete run
of the generation loop from the same seeded random grid.  With N GPUs the global grid is
(R*16384) x (C*16384) on an R x C domain grid, one process per GPU (weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line on rank 0.  `value` is device-timed (CUDA events around each step, L2
flushed between steps, max over ranks); `e2e` times the public API call including the H2D copy
of the packed initial grid from pinned memory and the D2H read of the result, every step.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PER_GPU = 16384
GENS = 1000
FREQ = 3
SEED = 20240229
METRIC = "cell-updates/sec (device-timed, max over ranks)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=PER_GPU, help="cells per side per GPU")
    ap.add_argument("--gens", type=int, default=GENS)
    ap.add_argument("--k", type=int, default=0, help="temporal depth override (0 = engine default)")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--no-similarity", action="store_true")
    return ap.parse_args()


def baseline_number():
    """BASELINE.md publishes no number for this metric (the reference has none)."""
    try:
        with open(os.path.join(ROOT, "BASELINE.json")) as f:
            pub = json.load(f).get("published") or {}
        v = pub.get("cell_updates_per_sec")
        return float(v) if v else None
    except Exception:
        return None


# ------------------------------------------------------------------ our arm
def bench_ours(args):
    import torch
    import torch.distributed as dist

    import gol_b200
    from gol_b200.ops import cpu
    from gol_b200.parallel import auto_grid
    from gol_b200.utils import ClockSampler

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    R, C = auto_grid(world)
    W, H = C * args.size, R * args.size
    freq = 0 if args.no_similarity else FREQ
    life = gol_b200.GameOfLife(W, H, gen_limit=args.gens, similarity_frequency=freq, grid=(R, C),
                               k_max=args.k, batch_gens=args.batch, device=local_rank)

    # the step's input: the packed initial grid in pinned host memory (every rank holds the
    # rows/columns it needs; for simplicity each rank keeps the whole global grid pinned)
    host = torch.from_numpy(cpu.random_grid(W, H, SEED).view("int64")).pin_memory()
    h2d_bytes = sum(((life.domain_extent(g)[2] + 63) // 64) * 8 * life.domain_extent(g)[3] for g in life.local_domains)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def one_step():
        life.stage_packed(host)         # H2D from pinned memory
        life.commit_staged()            # + first halo exchange
        return life.run()               # generation loop; reads back flags, generations, population

    sampler = ClockSampler(local_rank).start() if rank == 0 else None   # NVML set-up happens here, not in the timed region
    for _ in range(args.warmup):
        flush.fill_(1)
        res = one_step()
    barrier()
    if sampler:
        sampler.samples.clear()         # keep only the samples of the timed region

    launches0 = life.launch_count

    def timed(steps, do_flush=True, overlap=True, sync_ranks=False):
        """One timed region for all `steps` steps, run the way a serving loop would: the upload of
        step i+1's input (pinned host -> device, second stream) overlaps the generation loop of step i.
        Every step still uploads its own input and reads its own result back; the wall clock covers
        all of it (and the L2 flushes).  The device time is the CUDA-event time of each generation loop."""
        dev_ms, gens_done, gens_reported = 0.0, 0, 0
        barrier()
        t0 = time.perf_counter()
        life.stage_packed(host)
        for i in range(steps):
            if do_flush:
                flush.fill_(1)              # L2 flush between timed iterations
            torch.cuda.synchronize()
            # staged grid -> padded layout, first halo exchange; next step's H2D starts here and
            # overlaps this step's loop
            life.commit_staged(stage_next=host if overlap and i + 1 < steps else None)
            if sync_ranks:
                barrier()
            res = life.run()                # generation loop; reads back flags, generations, population
            if not overlap and i + 1 < steps:
                life.stage_packed(host)
            dev_ms += res.loop_ms           # CUDA events around the generation loop on the engine stream
            gens_done += res.generations_computed
            gens_reported = res.generations
            if os.environ.get("GOL_BENCH_DEBUG"):
                print(f"step {i}: wall so far {(time.perf_counter() - t0) * 1e3:.2f} ms, loop {res.loop_ms:.2f}", file=sys.stderr)
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3
        barrier()
        return dev_ms, e2e_ms, gens_done, gens_reported

    dev_ms, e2e_ms, gens_done, gens_reported = timed(args.steps)
    launches = life.launch_count - launches0
    clocks = sampler.stop() if sampler else None

    if os.environ.get("GOL_BENCH_VARIANTS"):   # diagnostics only: which part of a step costs what (stderr)
        for name, kw in (("default", {}), ("no-flush", {"do_flush": False}), ("serial-stage", {"overlap": False}),
                         ("rank-barrier-before-run", {"sync_ranks": True}),
                         ("no-flush,serial,barrier", {"do_flush": False, "overlap": False, "sync_ranks": True})):
            d, w, _, _ = timed(args.steps, **kw)
            tv = torch.tensor([d, w], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tv, op=dist.ReduceOp.MAX)
            if rank == 0:
                print(f"[variant] {name}: device {tv[0].item() / args.steps:.3f} ms/step, wall {tv[1].item() / args.steps:.3f} ms/step", file=sys.stderr)
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = t.tolist()
    cells = float(W) * float(H)
    value = cells * gens_done / (dev_ms * 1e-3)
    e2e = cells * gens_done / (e2e_ms * 1e-3)
    n_batches = -(-args.gens // (args.batch or 128)) + 1
    out = {
        "metric": METRIC, "value": value, "unit": "cell-updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / baseline_number()) if baseline_number() else None,
        "dtype": "1-bit cells (uint64 bit-planes; no floating point in this workload)",
        "data": "synthetic (seeded i.i.d. P(alive)=1/2 grid, generate.sh distribution)",
        "impl": "ours",
        "config": {"model": "Conway B3/S23 torus, reference generation loop", "grid": [W, H],
                   "cells_per_gpu": args.size * args.size, "generations": args.gens,
                   "generations_reported": gens_reported, "similarity_frequency": freq,
                   "global_batch": 1, "seq_len": None, "parallelism": f"2D domain decomposition {R}x{C}",
                   "k_max": args.k or "auto", "fast_path": life.uses_fast_path,
                   "l2": "256 MiB buffer written between timed steps (L2 flush); the loop itself is L2-resident by design"},
        "e2e": {"value": e2e, "unit": "cell-updates/s", "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(n_batches * 12 + 8)},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------ reference arm
def bench_reference(args):
    """Runs the UNMODIFIED reference CUDA program (src/game_cuda.cu) from baseline/_ref through
    its own CLI on the same grid; the metric is its own 'Execution time' line."""
    def unavailable(why):
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)

    if int(os.environ.get("RANK", "0")) != 0:
        return
    if args.gpus != 1:
        return unavailable("the reference CUDA program is single-GPU (no MPI+CUDA path) and its MPI programs "
                           "are CPU-only; no N>1 GPU reference exists")
    from gol_b200.ops import cpu
    from gol_b200.utils import ClockSampler, reference
    exes = {}
    for native in (False, True):
        try:
            exe = reference.build_cuda(arch_native=native)
        except Exception as e:  # noqa: BLE001
            exe = None
            err = str(e)
        if exe and os.path.exists(exe):
            exes["sm_100" if native else "as_shipped"] = exe
    if not exes:
        return unavailable("baseline/_ref has no reference build (sources unreachable or nvcc failed)")
    n = args.size
    if (n + 2) * (n + 2) >= 2 ** 31:
        return unavailable("reference int overflow for this size (defect D6)")
    work = tempfile.mkdtemp(prefix="golref_")
    inp = os.path.join(work, "input.txt")
    cpu.random_text_file(inp, n, n, SEED)
    results = {}
    sampler = ClockSampler(0).start()
    for label, exe in exes.items():
        # every run is a fresh process (text read, 1000 generations, text write: ~15 s), so warm-up only
        # matters for the PTX JIT cache of the as-shipped build: one warm-up run each; the as-shipped
        # build is timed once for the record, the sm_100 build for the requested number of steps
        n_warm = min(args.warmup, 1)
        n_timed = args.steps if label == "sm_100" or len(exes) == 1 else 1
        loop_ms, wall_ms, gens = [], [], None
        for i in range(n_warm + n_timed):
            t0 = time.perf_counter()
            g, ms, _, _ = reference.run_program(exe, n, n, inp, work, "cuda_output.out", timeout=3600)
            wall = (time.perf_counter() - t0) * 1e3
            if i >= n_warm:
                loop_ms.append(ms)
                wall_ms.append(wall)
                gens = g
        results[label] = {"loop_ms_per_step": sum(loop_ms) / len(loop_ms), "wall_ms_per_step": sum(wall_ms) / len(wall_ms),
                          "generations": gens, "timed_runs": n_timed}
    clocks = sampler.stop()
    best = min(results, key=lambda k: results[k]["loop_ms_per_step"])
    r = results[best]
    cells = float(n) * n
    gens = r["generations"] or args.gens
    value = cells * gens / (r["loop_ms_per_step"] * 1e-3)
    out = {
        "metric": METRIC, "value": value, "unit": "cell-updates/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["loop_ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "1 byte per cell (reference)", "data": "synthetic (same seeded grid, text file)",
        "impl": "reference",
        "config": {"model": "reference src/game_cuda.cu, unmodified, own CLI", "grid": [n, n], "generations": gens,
                   "build": best, "all_builds": results,
                   "timing": "the program's own 'Execution time' line (gettimeofday around its loop)"},
        "e2e": {"value": cells * gens / (r["wall_ms_per_step"] * 1e-3), "unit": "cell-updates/s",
                "note": "whole process wall time incl. its fgetc text read and fprintf write",
                "h2d_bytes_per_step": (n + 2) * (n + 2) * 2, "d2h_bytes_per_step": (n + 2) * (n + 2)},
        "gpu_launches": None,
        "clocks": clocks,
    }
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        bench_reference(a)
    else:
        bench_ours(a)
