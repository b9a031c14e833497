#!/usr/bin/env python
"""Headline benchmark: cell-updates/sec of the reference generation loop.

Config (BASELINE.json #2): 16384 x 16384 random torus per GPU, GEN_LIMIT = 1000 generations,
similarity check every 3 generations (the reference defaults).  One *step* is one complete run
of the generation loop from the same seeded random grid.  With N GPUs the global grid is
(R*16384) x (C*16384) on an R x C domain grid, one process per GPU (weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line on rank 0.
  value     device-timed (CUDA events around each step's generation loop, L2 flushed between
            steps, max over ranks).
  e2e       the same steps through the public API, wall clock: every step uploads its input
            (the rank's own packed blocks, pinned host memory), runs the loop and brings the
            FINAL GRID back into pinned host memory together with the result record.
  cli_wall  (N = 1) whole-process wall time of the a.out-compatible CLI on a text file: read +
            parse, generation loop, format + write - the same work the reference program does.
  verify    checksum / population of the final grid against an independent decomposition of
            the same global grid run inside one process.

The reference arm (--impl reference) drives the UNMODIFIED reference program from baseline/_ref
through its own CLI.  It imports nothing from this repository's package and loads none of its
native code: the input file is written with numpy, the clocks are sampled with NVML.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import shutil
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

PER_GPU = 16384
GENS = 1000
FREQ = 3
SEED = 20240229
METRIC = "cell-updates/sec (device-timed, max over ranks)"
L2_NOTE = ("no L2 reuse across timed steps: ours writes a 256 MiB buffer (> 126 MB L2) between steps, the reference "
           "runs every step in a fresh process on two 268 MB byte grids")


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
    ap.add_argument("--grid", default="", help="RxC decomposition override (default: the engine's choice for N GPUs)")
    ap.add_argument("--no-similarity", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-cli", action="store_true")
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


def common_config(w, h, gens, freq, parallelism):
    """The keys both arms print, so that equal work shows as equal configs."""
    return {"model": "Conway B3/S23 on a torus, reference generation loop (GEN_LIMIT, early exits)",
            "grid": [w, h], "generations": gens, "similarity_frequency": freq, "global_batch": 1, "seq_len": None,
            "parallelism": parallelism, "l2": L2_NOTE}


# ------------------------------------------------------- stand-alone helpers
class NvmlSampler:
    """SM clock / throttle-reason samples from NVML on a background thread (recipe's clocks line)."""

    def __init__(self, gpu_index=0, period_s=0.005):
        self.gpu, self.period = gpu_index, period_s
        self.samples = []
        self._stop = threading.Event()
        self._thread = None

    def _poll(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        except Exception:
            return self._poll_smi()
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((float(sm), float(mx), pw, [n for n, b in bits.items() if r & b]))
            except Exception:
                pass
            self._stop.wait(self.period)

    def _poll_smi(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5)
                f = [x.strip() for x in out.stdout.strip().splitlines()[0].split(",")]
                self.samples.append((float(f[0]), float(f[1]), float(f[2]),
                                     [n for n, v in zip(names, f[3:7]) if v.lower().startswith("active")]))
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        self._thread = threading.Thread(target=self._poll, daemon=True)
        self._thread.start()
        return self

    def clear(self):
        self.samples.clear()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = sorted({r for s in self.samples for r in s[3]})
        return {"sm_mhz": statistics.median(s[0] for s in self.samples), "sm_max_mhz": max(s[1] for s in self.samples),
                "reasons": reasons, "samples": len(self.samples), "power_w_max": max(s[2] for s in self.samples)}


def write_seeded_text_file(path, w, h, seed):
    """The benchmark's input grid as reference-format text, written with numpy only.  Same cells as the
    package's generator: bit i of word q of row y = bit i of splitmix64-finalised (seed, q, y)."""
    import numpy as np
    nq = (w + 63) // 64
    q = np.arange(nq, dtype=np.uint64)[None, :]
    rows_per = max(1, (64 << 20) // (w + 1))
    with np.errstate(over="ignore"), open(path, "wb") as f:
        for y0 in range(0, h, rows_per):
            y = np.arange(y0, min(h, y0 + rows_per), dtype=np.uint64)[:, None]
            z = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (y * np.uint64(0x100000001B3) + q + np.uint64(1))
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            z = z ^ (z >> np.uint64(31))
            bits = np.unpackbits(z.view(np.uint8).reshape(len(y), nq * 8), axis=1, bitorder="little")[:, :w]
            text = np.empty((len(y), w + 1), dtype=np.uint8)
            text[:, :w] = bits + ord("0")
            text[:, w] = ord("\n")
            f.write(text.tobytes())


def timed_cli(cmd, cwd, runs, warm):
    """Wall-clock milliseconds of whole-process runs of `cmd` (mean of `runs` after `warm` warm-ups) + last stdout."""
    walls, out = [], ""
    for i in range(warm + runs):
        t0 = time.perf_counter()
        p = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, timeout=3600, check=True)
        if i >= warm:
            walls.append((time.perf_counter() - t0) * 1e3)
        out = p.stdout
    return sum(walls) / len(walls), out


# ------------------------------------------------------------------ our arm
def bench_ours(args):
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import torch.distributed as dist

    import gol_b200
    from gol_b200.ops import dense
    from gol_b200.parallel import auto_grid

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cdev = dev
    if world > 1:
        import datetime
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # a rank that dies must not leave the others waiting for ten minutes
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    R, C = (int(v) for v in args.grid.split("x")) if args.grid else auto_grid(world)
    W, H = C * args.size, R * args.size
    freq = 0 if args.no_similarity else FREQ
    life = gol_b200.GameOfLife(W, H, gen_limit=args.gens, similarity_frequency=freq, grid=(R, C),
                               k_max=args.k, batch_gens=args.batch, device=local_rank)

    # the step's input and output: the rank's OWN packed blocks in pinned host memory
    n_local = len(life.local_domains)
    host_in, host_out = [], []
    for i, g in enumerate(life.local_domains):
        x0, y0, w, h = life.domain_extent(g)
        host_in.append(dense.random_grid(w, h, SEED, device=dev, x0=x0, y0=y0).cpu().pin_memory())
        host_out.append(torch.empty(life.local_block_shape(i), dtype=torch.int64).pin_memory())
    h2d_bytes = sum(b.numel() * 8 for b in host_in)
    d2h_bytes = sum(b.numel() * 8 for b in host_out)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    sampler = NvmlSampler(local_rank).start() if rank == 0 else None   # NVML set-up happens here, not in the timed region
    for _ in range(args.warmup):
        flush.fill_(1)
        life.stage_local(host_in)
        life.commit_staged()
        res = life.run()
        life.download_local(host_out)
    barrier()
    if sampler:
        sampler.clear()                 # keep only the samples of the timed region
    launches0 = life.launch_count

    def timed(steps):
        """One timed region for all `steps` steps, run the way a serving loop would: the upload of
        step i+1's input (pinned host -> device, second stream) overlaps the generation loop of step
        i; the download of step i's final grid is synchronous.  Every step uploads its own input and
        reads its own result back; the wall clock covers all of it (and the L2 flushes).  The device
        time is the CUDA-event time of each generation loop."""
        dev_ms, gens_done, res = 0.0, 0, None
        barrier()
        t0 = time.perf_counter()
        life.stage_local(host_in)
        for i in range(steps):
            flush.fill_(1)                  # L2 flush between timed iterations
            torch.cuda.synchronize()
            # staged grid -> padded layout, first halo exchange; the next step's H2D starts here
            life.commit_staged(stage_next_local=host_in if i + 1 < steps else None)
            res = life.run()                # generation loop; reads back flags, generations, population
            life.download_local(host_out)   # final grid -> pinned host memory
            dev_ms += res.loop_ms           # CUDA events around the generation loop on the engine stream
            gens_done += res.generations_computed
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3
        barrier()
        return dev_ms, e2e_ms, gens_done, res

    dev_ms, e2e_ms, gens_done, res = timed(args.steps)
    launches = life.launch_count - launches0
    clocks = sampler.stop() if sampler else None

    # identity of the final grid: decomposition-independent checksum + population, summed over the ranks
    lh = life.local_hash()
    ident = torch.tensor([lh - (1 << 64) if lh >= (1 << 63) else lh, life.population()], dtype=torch.int64, device=cdev)
    # the downloaded blocks are that grid: their host-side checksum must agree with the device's
    host_pop = int(sum(int(np.unpackbits(b.numpy().view(np.uint8)).sum()) for b in host_out))
    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=cdev)
    hp = torch.tensor([host_pop], dtype=torch.int64, device=cdev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(ident, op=dist.ReduceOp.SUM)
        dist.all_reduce(hp, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms = t.tolist()
    grid_hash, population = int(ident[0].item()) & ((1 << 64) - 1), int(ident[1].item())
    chunk_rows, deep_halo, fast_path = life.chunk_rows, life.uses_deep_halo, life.uses_fast_path

    verify = None
    if not args.no_verify:
        # the same global grid in ONE process on rank 0's GPU, cut differently (2x2 domains for a single
        # GPU, one domain per rank otherwise - borders then travel through local pushes only)
        life.close()
        if rank == 0:
            vgrid = (2, 2) if world == 1 else (R, C)
            v = gol_b200.GameOfLife(W, H, gen_limit=args.gens, similarity_frequency=freq, grid=vgrid, k_max=args.k,
                                    batch_gens=args.batch, device=local_rank, distributed=False)
            v.randomize(SEED)
            vres = v.run()
            verify = {"what": f"same {W}x{H} grid as {vgrid[0]}x{vgrid[1]} domains inside one process",
                      "hash_equal": v.local_hash() == grid_hash, "population_equal": v.population() == population,
                      "generations_equal": vres.generations == res.generations,
                      "downloaded_grid_population_equal": int(hp.item()) == population}
            v.close()
        barrier()

    cli = None
    if world == 1 and rank == 0 and not args.no_cli and args.size <= 32768:
        # like-for-like with the reference program's whole process: text in, text out
        work = tempfile.mkdtemp(prefix="golcli_")
        try:
            inp = os.path.join(work, "input.txt")
            write_seeded_text_file(inp, W, H, SEED)
            cmd = [os.path.join(ROOT, "bin", "gol"), str(W), str(H), inp, "--variant", "cuda", "--gen-limit", str(args.gens),
                   "--similarity-frequency", str(freq)]
            ms, out = timed_cli(cmd, work, runs=min(args.steps, 3), warm=1)
            g = re.search(r"Generations:\s+(\d+)", out)
            cli = {"ms": ms, "cmd": "bin/gol W H input.txt --variant cuda", "generations": int(g.group(1)) if g else None,
                   "what": "whole-process wall time: read + parse of the text grid, generation loop, format + write of "
                           "cuda_output.out", "input_bytes": os.path.getsize(inp),
                   "output_bytes": os.path.getsize(os.path.join(work, "cuda_output.out"))}
        except Exception as e:  # noqa: BLE001
            cli = {"error": str(e)[:300]}
        finally:
            shutil.rmtree(work, ignore_errors=True)

    cells = float(W) * float(H)
    value = cells * gens_done / (dev_ms * 1e-3)
    e2e = cells * gens_done / (e2e_ms * 1e-3)
    out = {
        "metric": METRIC, "value": value, "unit": "cell-updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / baseline_number()) if baseline_number() else None,
        "dtype": "1-bit cells (uint64 bit-planes; no floating point in this workload)",
        "data": "synthetic (seeded i.i.d. P(alive)=1/2 grid, generate.sh distribution)",
        "impl": "ours",
        "config": common_config(W, H, args.gens, freq, f"{R}x{C}"),
        "details": {"cells_per_gpu": args.size * args.size, "generations_reported": res.generations,
                    "exit_reason": res.exit_reason, "k_max": args.k or "auto", "fast_path": fast_path,
                    "chunk_rows": chunk_rows, "deep_halo_pairs": deep_halo, "decomposition": f"{R}x{C}",
                    "population": population, "grid_hash": f"{grid_hash:016x}"},
        "e2e": {"value": e2e, "unit": "cell-updates/s", "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes + 48),
                "what": "per step and rank: upload of the rank's packed blocks from pinned memory (overlaps the previous "
                        "step's loop), generation loop, download of the final grid into pinned memory + result record"},
        "cli_wall": cli,
        "verify": verify,
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------ reference arm
REF_SRC = os.environ.get("GOL_REFERENCE_DIR", "/root/reference")
REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def stage_reference():
    """baseline/_ref = verbatim copy of the reference checkout (idempotent); None if unreachable."""
    if os.path.exists(os.path.join(REF_DIR, "src", "game_cuda.cu")):
        return REF_DIR
    if not os.path.exists(os.path.join(REF_SRC, "src", "game_cuda.cu")):
        return None
    os.makedirs(REF_DIR, exist_ok=True)
    for sub in ("src", "include"):
        shutil.copytree(os.path.join(REF_SRC, sub), os.path.join(REF_DIR, sub), dirs_exist_ok=True)
    for f in ("Makefile", "generate.sh", "README.md"):
        shutil.copy(os.path.join(REF_SRC, f), os.path.join(REF_DIR, f))
    for root, dirs, files in os.walk(REF_DIR):   # the checkout is read-only; the copy must not be
        for name in dirs + files:
            path = os.path.join(root, name)
            os.chmod(path, os.stat(path).st_mode | 0o200)
    return REF_DIR


def build_reference_cuda(native):
    """`nvcc src/game_cuda.cu` as shipped (reference Makefile:27-28; PTX-JIT on B200) or with -O3 -arch=sm_100."""
    if stage_reference() is None or shutil.which("nvcc") is None:
        return None
    exe = os.path.join(REF_DIR, "game_cuda_sm100" if native else "game_cuda_asshipped")
    if not os.path.exists(exe):
        cmd = ["nvcc"] + (["-O3", "-arch=sm_100"] if native else []) + ["src/game_cuda.cu", "-o", exe]
        subprocess.run(cmd, cwd=REF_DIR, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return exe


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
    exes = {}
    for native in (False, True):
        try:
            exe = build_reference_cuda(native)
        except Exception:  # noqa: BLE001
            exe = None
        if exe and os.path.exists(exe):
            exes["sm_100" if native else "as_shipped"] = exe
    if not exes:
        return unavailable("baseline/_ref has no reference build (sources unreachable or nvcc failed)")
    n = args.size
    if (n + 2) * (n + 2) >= 2 ** 31:
        return unavailable("reference int overflow for this size (defect D6)")
    if args.gens != GENS or args.no_similarity:
        return unavailable("the reference's GEN_LIMIT and similarity settings are compile-time constants (1000, every 3)")
    work = tempfile.mkdtemp(prefix="golref_")
    inp = os.path.join(work, "input.txt")
    write_seeded_text_file(inp, n, n, SEED)
    results = {}
    sampler = NvmlSampler(0).start()
    for label, exe in exes.items():
        # every run is a fresh process (text read, 1000 generations, text write: ~15 s), so warm-up only
        # matters for the PTX JIT cache of the as-shipped build: one warm-up run each; the as-shipped
        # build is timed once for the record, the sm_100 build for the requested number of steps
        n_warm = min(args.warmup, 1)
        n_timed = args.steps if label == "sm_100" or len(exes) == 1 else 1
        loop_ms, wall_ms, gens = [], [], None
        for i in range(n_warm + n_timed):
            t0 = time.perf_counter()
            p = subprocess.run([exe, str(n), str(n), inp], cwd=work, capture_output=True, text=True, timeout=3600, check=True)
            wall = (time.perf_counter() - t0) * 1e3
            g = re.search(r"Generations:\s+(\d+)", p.stdout)
            ms = re.search(r"Execution time:\s+([0-9.]+)", p.stdout)
            if i >= n_warm:
                loop_ms.append(float(ms.group(1)))
                wall_ms.append(wall)
                gens = int(g.group(1)) if g else None
        results[label] = {"loop_ms_per_step": sum(loop_ms) / len(loop_ms), "wall_ms_per_step": sum(wall_ms) / len(wall_ms),
                          "generations": gens, "timed_runs": n_timed}
    clocks = sampler.stop()
    out_bytes = os.path.getsize(os.path.join(work, "cuda_output.out")) if os.path.exists(os.path.join(work, "cuda_output.out")) else None
    in_bytes = os.path.getsize(inp)
    shutil.rmtree(work, ignore_errors=True)
    best = min(results, key=lambda k: results[k]["loop_ms_per_step"])
    r = results[best]
    cells = float(n) * n
    gens = r["generations"] or args.gens
    value = cells * gens / (r["loop_ms_per_step"] * 1e-3)
    out = {
        "metric": METRIC, "value": value, "unit": "cell-updates/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["loop_ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "1 byte per cell (reference)",
        "data": "synthetic (seeded i.i.d. P(alive)=1/2 grid, generate.sh distribution)",
        "impl": "reference",
        "config": common_config(n, n, gens, FREQ, "1x1"),
        "details": {"program": "reference src/game_cuda.cu, unmodified, own CLI", "build": best, "all_builds": results,
                    "timing": "the program's own 'Execution time' line (gettimeofday around its loop)"},
        "e2e": {"value": cells * gens / (r["wall_ms_per_step"] * 1e-3), "unit": "cell-updates/s",
                "ms_per_step": r["wall_ms_per_step"],
                "what": "whole process wall time incl. its fgetc text read and fprintf write (compare with cli_wall of our arm)",
                "h2d_bytes_per_step": (n + 2) * (n + 2) * 2, "d2h_bytes_per_step": (n + 2) * (n + 2)},
        "cli_wall": {"ms": r["wall_ms_per_step"], "cmd": "a.out W H input.txt", "generations": gens,
                     "what": "whole-process wall time: read + parse of the text grid, generation loop, write of cuda_output.out",
                     "input_bytes": in_bytes, "output_bytes": out_bytes},
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
