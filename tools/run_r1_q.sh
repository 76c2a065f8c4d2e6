#!/bin/bash
# pass Q: same-box A/B of the c3 step with the cluster-multicast GEMM (VC_GEMM_MC=1) against the plain one.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; }
for i in 1 2; do
  VC_GEMM_MC=1 T=150 run q_mc_$i python bench.py --steps 3 --warmup 3 --no-cpu-baseline
  T=150 run q_plain_$i python bench.py --steps 3 --warmup 3 --no-cpu-baseline
done
