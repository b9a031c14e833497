#!/bin/bash
# 8-GPU measurement suite (one box): bench curve, decomposition / deep-halo / PDL matrix at 8 GPUs, strong
# scaling at 65536^2, per-launch trace.  Results under gpurun_out/.  EVERY multi-rank command runs under its
# own timeout: eight idle GPUs behind a hung collective are charged like eight busy ones.
OUT=gpurun_out; mkdir -p $OUT
tr() { timeout ${TMO:-150} python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "${@:2}"; }
mg() { tr $1 scripts/mgpu_check.py "${@:2}" 2>&1 | grep -E "^\[(time|check)\]|Error|error|Traceback"; }
for n in 8 4 2; do
  echo "== bench N=$n"; tr $n bench.py --gpus $n --steps 5 --warmup 3 > $OUT/bench_n$n.json 2> $OUT/bench_n$n.err || echo "bench N=$n failed or timed out"
  tail -c 1500 $OUT/bench_n$n.json; grep -v "^\*\|OMP_NUM" $OUT/bench_n$n.err | tail -3 | cut -c1-300
done
echo "== correctness 8x1"; mg 8 --grid 8x1 --size 2048 --gens 64
for cfg in "8x1 1 2" "8x1 0 1" "2x4 1 2"; do
  set -- $cfg
  echo "== weak 16384^2/GPU grid $1 GOL_DEEP_HALO=$2 GOL_PDL=$3"
  GOL_DEEP_HALO=$2 GOL_PDL=$3 mg 8 --skip-check --grid $1 --size 16384
done
for cfg in "8x1 1 2" "2x4 1 2"; do
  set -- $cfg
  echo "== strong 65536^2 grid $1 GOL_DEEP_HALO=$2 GOL_PDL=$3"
  GOL_DEEP_HALO=$2 GOL_PDL=$3 mg 8 --skip-check --grid $1 --width 65536 --height 65536
done
echo "== 1 GPU 65536^2"; mg 1 --skip-check --width 65536 --height 65536
echo "== trace 8x1"; tr 8 scripts/trace_run.py --grid 8x1 2>&1 | grep "^\[rank [07]\]"
