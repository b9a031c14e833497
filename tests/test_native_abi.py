"""Builds a plain-C client of include/gol_b200.h with gcc, links it against the in-tree library and
runs it: the C ABI is usable without C++ or Python (the reference is a C code base)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def exe(tmp_path_factory, lib):
    out = tmp_path_factory.mktemp("native") / "c_abi_smoke"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "native", "c_abi_smoke.c"), "-o", str(out),
                    "-L", ROOT, "-lgolb200", f"-Wl,-rpath,{ROOT}"], check=True)
    return str(out)


def test_c_abi_cpu(exe):
    p = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert "c abi ok" in p.stdout


@pytest.mark.gpu
def test_c_abi_gpu(exe, cuda_device):
    p = subprocess.run([exe, "gpu"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert "gpu: generations=50" in p.stdout and "fast_path=1" in p.stdout
