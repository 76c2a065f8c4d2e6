#!/bin/bash
# pass K (1 GPU, one box): GEMM rasterisation group height (A-panel L2 budget 22 / 44 / 64 MB): DRAM traffic per launch
# (ncu) and the whole c3 step.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -2 gpurun_out/$name.log | cut -c1-300; }
for mb in 22 44 64; do
  export VC_GEMM_PANEL_MB=$mb
  T=120 run k_traffic_$mb ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_bf16_kernel --csv --log-file gpurun_out/k_traffic_$mb.csv python tools/gemm_traffic.py
done
for mb in 22 44 64 22 44; do
  export VC_GEMM_PANEL_MB=$mb
  T=200 run k_bench_${mb}_$RANDOM python bench.py --steps 3 --warmup 3 --no-cpu-baseline
done
