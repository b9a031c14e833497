#!/bin/bash
# CPU baselines of the reference on the GPU box's host (configs 1, 3, 4 of BASELINE.json).
# Config 1 runs the unmodified serial program.  The MPI programs at 32768^2 / 65536^2 would need
# hours for GEN_LIMIT = 1000 (about 10 s per generation), so for TIMING ONLY they are rebuilt from a
# copy whose one line `#define GEN_LIMIT 1000` reads `#define GEN_LIMIT <G>` (and, for config 4,
# without `#define CHECK_SIMILARITY`, as README.md:65 of the reference prescribes); everything else is
# the reference source, built by the reference Makefile against baseline/mpishim.
set -e
cd "$(dirname "$0")/.."
ROOT=$PWD
OUT=$ROOT/gpurun_out/ref_cpu
mkdir -p $OUT
W=/tmp/refcpu; rm -rf $W; mkdir -p $W; cd $W
nproc > $OUT/nproc.txt
# ---- config 1: serial, 840 x 840, GEN_LIMIT 1000, similarity every 3
$ROOT/bin/gol-generate 840 840 7 > in840.txt
for i in 1 2 3; do $ROOT/baseline/_ref/game_serial 840 840 in840.txt | grep -E "Generations|Execution"; done > $OUT/config1_serial.txt
cp game_output.out ref840.out
$ROOT/bin/gol 840 840 in840.txt --stats > $OUT/config1_ours.txt 2>&1; cmp game_output.out ref840.out && echo "ours == reference output" >> $OUT/config1_ours.txt
# ---- timing copies with a small GEN_LIMIT
G=${GENS:-4}   # must be even: after an odd number of generations the reference reads a freed buffer (defect D12) and segfaults at these sizes
T=$W/ref_timing; mkdir -p $T; cp -r $ROOT/baseline/_ref/src $ROOT/baseline/_ref/include $ROOT/baseline/_ref/Makefile $T/
sed -i "s/^#define GEN_LIMIT 1000/#define GEN_LIMIT $G/" $T/src/*.c
( cd $T && for t in collective openmp; do make $t MPICC=$ROOT/baseline/mpishim/mpicc >/dev/null 2>&1 && mv a.out game_$t; done )
T2=$W/ref_timing_nosim; mkdir -p $T2; cp -r $T/src $T/include $T/Makefile $T2/
sed -i "s/^#define CHECK_SIMILARITY//" $T2/src/*.c
( cd $T2 && for t in collective openmp; do make $t MPICC=$ROOT/baseline/mpishim/mpicc >/dev/null 2>&1 && mv a.out game_$t; done )
MPIEXEC=$ROOT/baseline/mpishim/mpiexec
# ---- config 3: 65536 x 65536, 4 ranks (the only torus-correct multi-rank count), similarity on
if [ "${BIG:-1}" = 1 ]; then
$ROOT/bin/gol-generate 65536 65536 3 -o in65536.txt
for t in collective openmp; do
  $MPIEXEC -n 4 $T/game_$t 65536 65536 in65536.txt > $OUT/config3_${t}_n4_gen$G.txt 2>&1 || echo "FAILED rc=$?" >> $OUT/config3_${t}_n4_gen$G.txt
  rm -f *_output.out
done
rm -f in65536.txt
fi
# ---- config 4: weak scaling, 32768^2 per rank, similarity off: 1 rank and 4 ranks
$ROOT/bin/gol-generate 32768 32768 4 -o in32768.txt
$MPIEXEC -n 1 $T2/game_collective 32768 32768 in32768.txt > $OUT/config4_collective_n1_gen$G.txt 2>&1 || true
rm -f *_output.out in32768.txt
if [ "${BIG:-1}" = 1 ]; then
$ROOT/bin/gol-generate 65536 65536 4 -o in65536b.txt
for t in collective openmp; do $MPIEXEC -n 4 $T2/game_$t 65536 65536 in65536b.txt > $OUT/config4_${t}_n4_gen$G.txt 2>&1 || true; rm -f *_output.out; done
fi
cd /; rm -rf $W
grep -H . $OUT/*.txt | grep -vE "time.txt" | head -60
