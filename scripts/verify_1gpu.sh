#!/bin/bash
# One-GPU verification pass of the final tree: GPU tests, bench line (ours), per-launch trace of run().
OUT=gpurun_out; mkdir -p $OUT
echo "== pytest"; timeout 420 python -m pytest tests -m gpu -q --timeout 150 2>&1 | tail -12 > $OUT/pytest_gpu.log; cat $OUT/pytest_gpu.log
echo "== bench"; timeout 200 python bench.py --steps 5 --warmup 3 > $OUT/bench_n1.json 2> $OUT/bench_n1.err; cat $OUT/bench_n1.json | cut -c1-2600; tail -3 $OUT/bench_n1.err
echo "== trace"; timeout 120 python scripts/trace_run.py --out $OUT/trace_run 2>&1 | tail -4
timeout 60 python scripts/trace_run.py --out $OUT/trace_step --step 256 2>&1 | tail -2
