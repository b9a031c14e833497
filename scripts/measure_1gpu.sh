#!/bin/bash
# One-GPU measurement pass (scripts/measure_1gpu.sh): GPU test suite, tile-kernel experiment, CLI phase times, early-exit cost,
# per-launch trace of run().  Every step under its own timeout.
OUT=gpurun_out; mkdir -p $OUT
echo "== pytest"; timeout 420 python -m pytest tests -m gpu -q --timeout 150 2>&1 | tail -25 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
echo "== tilebench"; timeout 300 scripts/tilebench.sh 16384 16384 20 > $OUT/tilebench_16384.txt 2>&1; cat $OUT/tilebench_16384.txt
echo "== cli"
timeout 300 python - > $OUT/cli_phases.txt 2>&1 <<'PY'
import os, subprocess, tempfile, time
root = os.getcwd()
work = tempfile.mkdtemp(prefix="golcli_")
subprocess.run([root + "/bin/gol-generate", "16384", "16384", "20240229", "-o", work + "/in.txt"], check=True, timeout=120)
for variant in ("collective", "cuda"):
    for i in range(2):
        t0 = time.perf_counter()
        p = subprocess.run([root + "/bin/gol", "16384", "16384", "in.txt", "--variant", variant, "--stats"], cwd=work,
                           capture_output=True, text=True, timeout=120)
        wall = time.perf_counter() - t0
        print(f"[cli] --variant {variant} run {i}: wall {wall * 1e3:.0f} ms | " + " | ".join((p.stdout + p.stderr).split("\n")), flush=True)
PY
cat $OUT/cli_phases.txt
echo "== early exit"; timeout 120 python scripts/early_exit_timing.py > $OUT/early_exit.txt 2>&1; cat $OUT/early_exit.txt
echo "== trace"; timeout 120 python scripts/trace_run.py --out $OUT/trace_run 2>&1 | tail -4
