#!/bin/bash
# 1 -> 8 GPU scaling run on one box: correctness at 8 ranks, weak-scaling bench lines, 65536^2 strong config.
set -x
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29512"
mkdir -p gpurun_out
nvidia-smi topo -m | head -12 > gpurun_out/topo.txt
timeout 600 $T --nproc-per-node 8 scripts/mgpu_check.py 2>&1 | grep -E "check|time|rror" | tee gpurun_out/mgpu8.log
for n in 1 2 4 8; do
  if [ $n = 1 ]; then timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 | tee gpurun_out/bench_n$n.json
  else timeout 600 $T --nproc-per-node $n bench.py --gpus $n --steps 5 --warmup 3 2>/dev/null | grep '^{' | tee gpurun_out/bench_n$n.json; fi
done
# config 3: 65536 x 65536 on 2x4, reference loop (similarity every 3)
timeout 600 $T --nproc-per-node 8 scripts/mgpu_check.py --skip-check --width 65536 --height 65536 --gens 256 2>&1 | grep -E "time|rror" | tee gpurun_out/strong65536.log
# same grid on one GPU for the strong-scaling efficiency
timeout 600 python scripts/profile_step.py --size 65536 --gens 256 --reps 2 | tee gpurun_out/single65536.log
