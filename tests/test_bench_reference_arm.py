"""The reference arm of bench.py must stand on its own: nothing of this repository's package or native
library may be loaded by its process (the driver checks the process map), and the input file it writes
with numpy must be the very grid the package's generator produces for our arm."""
import ast
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.argv = argv
    return mod


def test_numpy_generator_matches_native_generator(tmp_path, lib):
    from gol_b200.ops import cpu
    bench = load_bench()
    for w, h, seed in [(200, 37, 5), (64, 3, 1), (1000, 17, 20240229), (129, 70, 7)]:
        a, b = tmp_path / "a.txt", tmp_path / "b.txt"
        bench.write_seeded_text_file(str(a), w, h, seed)
        cpu.random_text_file(str(b), w, h, seed)
        assert a.read_bytes() == b.read_bytes(), (w, h, seed)


def test_reference_arm_never_touches_the_package():
    tree = ast.parse(open(os.path.join(ROOT, "bench.py")).read())
    reference_side = {"bench_reference", "stage_reference", "build_reference_cuda", "write_seeded_text_file", "timed_cli",
                      "NvmlSampler", "common_config", "baseline_number", "parse_args"}
    for node in tree.body:
        name = getattr(node, "name", None)
        if name in reference_side:
            src = ast.get_source_segment(open(os.path.join(ROOT, "bench.py")).read(), node)
            assert "gol_b200" not in src and "libgolb200" not in src, name
    # module level: no import of the package either
    for node in tree.body:
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            names = [a.name for a in node.names] + [getattr(node, "module", "") or ""]
            assert not any("gol_b200" in n for n in names)


def test_reference_arm_process_loads_no_native_library():
    # N = 2 has no reference (the reference CUDA program is single-GPU): the arm answers "unavailable" and exits 0;
    # the process must get there without having mapped libgolb200.so
    code = ("import sys, runpy\n"
            "sys.argv = ['bench.py', '--impl', 'reference', '--gpus', '2']\n"
            f"runpy.run_path({os.path.join(ROOT, 'bench.py')!r}, run_name='__main__')\n"
            "maps = open('/proc/self/maps').read()\n"
            "assert 'libgolb200' not in maps, 'native library mapped in the reference arm'\n"
            "assert 'gol_b200' not in sys.modules\n")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][0]
    d = json.loads(line)
    assert d["impl"] == "reference" and "unavailable" in d
