#!/bin/bash
# Two-GPU pass: bench line, correctness incl. checksum against the single-process run, per-launch trace.
OUT=gpurun_out; mkdir -p $OUT
tr() { timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
echo "== bench N=2"; tr bench.py --gpus 2 --steps 5 --warmup 3 > $OUT/bench_n2.json 2> $OUT/bench_n2.err; tail -c 2500 $OUT/bench_n2.json; grep -v "^\*\|OMP_NUM" $OUT/bench_n2.err | tail -5 | cut -c1-300
echo "== correctness 2x1"; tr scripts/mgpu_check.py --grid 2x1 --size 2048 --gens 32 2>&1 | grep -E "^\[(time|check)\]|Error|error|Traceback" | cut -c1-300
echo "== trace 2x1"; tr scripts/trace_run.py --grid 2x1 --out $OUT/trace_run 2>&1 | grep "^\[rank" | cut -c1-300
