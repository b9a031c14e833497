"""CPU check of the strip kernel's work-item geometry: tests/native/items_check.cpp includes
csrc/life_items.h (the functions the kernels execute, compiled here for the host) and verifies for
thousands of domain shapes that the items tile the computed region exactly once, that only flagged
items can touch a ghost zone, that the flag counters' targets match and that border items are numbered
first (pushing launches) or last (waiting launches)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_item_geometry_invariants(tmp_path):
    exe = tmp_path / "items_check"
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "native", "items_check.cpp"), "-o", str(exe)], check=True)
    p = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "items ok" in p.stdout, p.stdout[-3000:]
