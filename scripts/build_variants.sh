#!/bin/bash
# Builds kernel-variant libraries for A/B timing: variants/libgol_s<SKEW>_o<ONE_SIDED>.so
# (run a variant with GOL_LIB=variants/libgol_s1_o1.so python scripts/profile_step.py ...)
set -e
cd "$(dirname "$0")/.."
mkdir -p variants build
ARCH="-gencode arch=compute_100a,code=sm_100a"
for s in 0 1; do for o in 0 1; do
  nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -DGOL_SKEW=$s -DGOL_ONE_SIDED=$o -c csrc/life_kernels.cu -o build/life_kernels_s${s}_o${o}.o &
done; done; wait
for s in 0 1; do for o in 0 1; do
  nvcc $ARCH -shared -o variants/libgol_s${s}_o${o}.so build/life_kernels_s${s}_o${o}.o build/dense_kernels.o build/engine.o build/static_scratch.o build/host_util.o
done; done
ls -la variants
