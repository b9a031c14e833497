#!/bin/bash
# One-GPU pass: tile-kernel experiment (one process per variant) and the ncu captures of every kernel.
OUT=gpurun_out; mkdir -p $OUT
echo "== tilebench"; timeout 300 scripts/tilebench.sh 16384 16384 20 > $OUT/tilebench_16384.txt 2>&1; cat $OUT/tilebench_16384.txt
echo "== ncu strip kernels (launches 13,14 = <8,false>, 15 = <8,true> of the first batch)"
GOL_CHUNK_ROWS=85 timeout 400 ncu --set full --clock-control none --import-source on -k regex:life_strip_kernel -s 13 -c 3 \
    -f -o $OUT/prof_strip python scripts/profile_run.py > $OUT/ncu_strip.log 2>&1; tail -4 $OUT/ncu_strip.log
echo "== ncu auxiliary kernels"
timeout 400 ncu --set full --clock-control none --import-source on \
    -k regex:'pack_kernel|unpack_kernel|halo_push|scan_alive|gather_flags|dense_to_padded|padded_to_dense|popcount|hash_padded|step_dense|random_dense' \
    -c 16 -f -o $OUT/prof_aux python scripts/io_case.py > $OUT/ncu_aux.log 2>&1; tail -4 $OUT/ncu_aux.log
ls -la $OUT/*.ncu-rep
