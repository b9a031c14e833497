"""Builds and runs the UNMODIFIED reference programs as oracle / baseline.

Sources are copied verbatim from the read-only reference checkout into ``baseline/_ref/src``
(git-ignored, but it travels to the GPU box with the snapshot) and compiled there with the
reference's own flags (``Makefile:10-13, 27-28`` of the reference).  Nothing from these
sources is used by the product.
"""
from __future__ import annotations

import os
import re
import shutil
import subprocess

from .._native import REPO_ROOT

REF_SRC = os.environ.get("GOL_REFERENCE_DIR", "/root/reference")
REF_DIR = os.path.join(REPO_ROOT, "baseline", "_ref")


def stage_sources() -> str | None:
    """Copy the reference tree into baseline/_ref (idempotent); None if it is not reachable."""
    dst = os.path.join(REF_DIR, "src")
    if os.path.exists(os.path.join(dst, "game.c")):
        return REF_DIR
    if not os.path.exists(os.path.join(REF_SRC, "src", "game.c")):
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


def build_serial() -> str | None:
    """gcc -std=c99 -Wall -O3 src/game.c (the reference's `make game`)."""
    if stage_sources() is None:
        return None
    exe = os.path.join(REF_DIR, "game_serial")
    if not os.path.exists(exe):
        subprocess.run(["gcc", "-std=c99", "-Wall", "-O3", "src/game.c", "-o", exe], cwd=REF_DIR, check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return exe


def build_cuda(arch_native: bool = False) -> str | None:
    """`nvcc src/game_cuda.cu` as shipped (PTX-JIT on B200) or with -O3 -arch=sm_100."""
    if stage_sources() is None or shutil.which("nvcc") is None:
        return None
    exe = os.path.join(REF_DIR, "game_cuda_sm100" if arch_native else "game_cuda_asshipped")
    if not os.path.exists(exe):
        cmd = ["nvcc"] + (["-O3", "-arch=sm_100"] if arch_native else []) + ["src/game_cuda.cu", "-o", exe]
        subprocess.run(cmd, cwd=REF_DIR, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return exe


def run_program(exe: str, width: int, height: int, input_path: str, workdir: str, output_name: str,
                timeout: float | None = None):
    """Runs `exe W H input` in workdir; returns (generations, execution_ms, output bytes, stdout)."""
    p = subprocess.run([exe, str(width), str(height), os.path.abspath(input_path)], cwd=workdir,
                       capture_output=True, text=True, timeout=timeout, check=True)
    gens = re.search(r"Generations:\s+(\d+)", p.stdout)
    ms = re.search(r"Execution time:\s+([0-9.]+)", p.stdout)
    with open(os.path.join(workdir, output_name), "rb") as f:
        out = f.read()
    return (int(gens.group(1)) if gens else None, float(ms.group(1)) if ms else None, out, p.stdout)


MPI_VARIANTS = {"mpi": "mpi_output.out", "collective": "collective_output.out", "async": "async_output.out",
                "openmp": "openmp_output.out"}


def build_mpi() -> dict | None:
    """The four MPI programs, unmodified, through the reference's own Makefile targets with the
    single-node shim's mpicc (baseline/mpishim).  Returns {variant: exe, "mpiexec": launcher}."""
    if stage_sources() is None:
        return None
    base = os.path.join(REPO_ROOT, "baseline")
    exes = {v: os.path.join(REF_DIR, f"game_{v}") for v in MPI_VARIANTS}
    launcher = os.path.join(base, "mpishim", "mpiexec")
    if not all(os.path.exists(p) for p in list(exes.values()) + [launcher]):
        subprocess.run(["make", "-C", base, "shim", "ref_mpi"], check=True, stdout=subprocess.DEVNULL,
                       stderr=subprocess.DEVNULL)
    exes["mpiexec"] = launcher
    return exes


def run_mpi(exes: dict, variant: str, ranks: int, width: int, input_path: str, workdir: str,
            timeout: float | None = None):
    """`mpiexec -n ranks ./a.out W W input` for an MPI variant under the shim; same return as run_program
    plus the 'Reading file' / 'Writing file' times in a dict."""
    p = subprocess.run([exes["mpiexec"], "-n", str(ranks), exes[variant], str(width), str(width),
                        os.path.abspath(input_path)], cwd=workdir, capture_output=True, text=True, timeout=timeout,
                       check=True)
    gens = re.search(r"Generations:\s+(\d+)", p.stdout)
    ms = re.search(r"Execution time:\s+([0-9.]+)", p.stdout)
    rd = re.search(r"Reading file:\s+([0-9.]+)", p.stdout)
    wr = re.search(r"Writing file:\s+([0-9.]+)", p.stdout)
    with open(os.path.join(workdir, MPI_VARIANTS[variant]), "rb") as f:
        out = f.read()
    return (int(gens.group(1)) if gens else None, float(ms.group(1)) if ms else None, out,
            {"read_ms": float(rd.group(1)) if rd else None, "write_ms": float(wr.group(1)) if wr else None,
             "stdout": p.stdout})
