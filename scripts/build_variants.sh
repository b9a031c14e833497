#!/bin/bash
# Builds the experiment binaries (not part of `make all`; nothing here is linked into libgolb200.so):
#   bin/gol-tilebench   TMA-fed cluster/DSMEM tile kernel against the product's register pipeline
#                       (csrc/experiments/life_tile.cu)
# The round-1 kernel variants (skewed levels, IMAD shifts, warp-pipelined FIFOs, persistent dataflow
# kernel) are kept for the record in csrc/experiments/life_kernels_r1_variants.cu; they were written
# against the round-1 engine interface and are no longer built.
set -e
cd "$(dirname "$0")/.."
make bin/gol-tilebench
grep -E "life_tile_kernel|registers|spill" build/life_tile.ptxas.log | grep -A1 "life_tile_kernel" | grep -E "registers" | sort | uniq -c
ls -la bin/gol-tilebench
