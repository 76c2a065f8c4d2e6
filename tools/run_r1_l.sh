#!/bin/bash
# pass L (1 GPU, one box): TMA L2 eviction hints (A evict_last, W evict_first, streaming C stores) x group height.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -1 gpurun_out/$name.log | cut -c1-200; }
for cfg in 3:16 3:32 3:48 7:48; do
  export VC_GEMM_L2_HINT=${cfg%%:*} VC_GEMM_GROUP_M=${cfg##*:}
  T=100 run l_traffic_h${VC_GEMM_L2_HINT}_g${VC_GEMM_GROUP_M} ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_bf16_kernel --csv --log-file gpurun_out/l_traffic_h${VC_GEMM_L2_HINT}_g${VC_GEMM_GROUP_M}.csv python tools/gemm_traffic.py
done
for cfg in 0:16 3:32 3:48 7:48; do
  export VC_GEMM_L2_HINT=${cfg%%:*} VC_GEMM_GROUP_M=${cfg##*:}
  T=150 run l_bench_h${VC_GEMM_L2_HINT}_g${VC_GEMM_GROUP_M} python bench.py --steps 3 --warmup 3 --no-cpu-baseline
done
