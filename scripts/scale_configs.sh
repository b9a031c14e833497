#!/bin/bash
# Configs 3, 4, 5 of BASELINE.json on an 8-GPU box (ours).  Device-timed, max over ranks.
set -x
NS=${NS:-"1 2 4 8"}; ONLY5=${ONLY5:-0}   # GPU counts to run (the box must have at least the largest)
MAXN=$(echo $NS | tr " " "\n" | sort -n | tail -1)
cd "$(dirname "$0")/.."
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29512"
OUT=gpurun_out/configs; mkdir -p $OUT
if [ $ONLY5 = 0 ]; then
# weak-scaling bench lines (config 2 per GPU)
for n in $NS; do
  if [ $n = 1 ]; then timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 > $OUT/bench_n$n.json
  else timeout 600 $T --nproc-per-node $n bench.py --gpus $n --steps 5 --warmup 3 2>/dev/null | grep '^{' > $OUT/bench_n$n.json; fi
done
# config 3: 65536^2 on 2x4, reference loop; and on one GPU for the strong-scaling efficiency
[ $MAXN = 8 ] && timeout 600 $T --nproc-per-node 8 scripts/mgpu_check.py --skip-check --width 65536 --height 65536 --gens 1024 2>&1 | grep -E "time|rror" > $OUT/config3_8gpu.txt
echo $NS | grep -qw 1 && timeout 600 python scripts/profile_step.py --size 65536 --gens 256 --reps 2 > $OUT/config3_1gpu.txt
# config 4: weak scaling 32768^2 per GPU, similarity off (pure stencil + halo = step())
for n in $NS; do
  if [ $n = 1 ]; then timeout 600 python scripts/profile_step.py --size 32768 --gens 512 --reps 2 > $OUT/config4_n$n.txt
  else timeout 600 $T --nproc-per-node $n scripts/mgpu_check.py --skip-check --size 32768 --gens 512 2>&1 | grep -E "time. step|rror" > $OUT/config4_n$n.txt; fi
done
fi
# config 5: strong scaling incl. parallel text read + collective-format write
S=${BIGSIZE:-131072}
if [ $MAXN = 8 ]; then
mkdir -p /tmp/cfg5 && bin/gol-generate $S $S 5 -o /tmp/cfg5/in.txt && ls -la /tmp/cfg5 > $OUT/config5_files.txt
( cd /tmp/cfg5 && $OLDPWD/bin/gol $S $S in.txt --variant collective --gpus 8 --stats > $OLDPWD/$OUT/config5_8gpu.txt 2>&1; ls -la >> $OLDPWD/$OUT/config5_files.txt; head -c 64 collective_output.out | head -1 | cut -c1-64 >> $OLDPWD/$OUT/config5_files.txt )
rm -rf /tmp/cfg5
fi
grep -H . $OUT/*.txt | cut -c1-300
