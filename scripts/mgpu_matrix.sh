#!/bin/bash
# Multi-GPU timing matrix (run on the GPU box): decomposition x deep-halo pairs x PDL across processes.
#   scripts/mgpu_matrix.sh <ngpus> "<grids>" [size] [extra mgpu_check args]
N=${1:-2}; GRIDS=${2:-"2x1 1x2"}; SIZE=${3:-16384}; shift 3
OUT=gpurun_out; mkdir -p $OUT
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) scripts/mgpu_check.py "${@:2}" 2>&1 | grep -E "^\[(time|check)\]|Error|error|Traceback" ; }
echo "== 1 GPU baseline"; run 1 --skip-check --size $SIZE "$@"
for g in $GRIDS; do
  echo "== $g correctness (defaults)"; run $N --grid $g --size 4096 --gens 64 "$@"
  for dh in 0 1; do for pdl in 1 2; do
    echo "== $g GOL_DEEP_HALO=$dh GOL_PDL=$pdl"
    GOL_DEEP_HALO=$dh GOL_PDL=$pdl run $N --skip-check --grid $g --size $SIZE "$@"
  done; done
done
