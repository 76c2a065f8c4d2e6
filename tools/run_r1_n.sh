#!/bin/bash
# pass N: same-box A/B of the GEMM rasterisation/hint default (new) against the previous setting (old).
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; }
for i in 1 2; do
  VC_GEMM_L2_HINT=0 VC_GEMM_GROUP_M=16 T=150 run n_old_$i python bench.py --steps 3 --warmup 3 --no-cpu-baseline
  T=150 run n_new_$i python bench.py --steps 3 --warmup 3 --no-cpu-baseline
done
