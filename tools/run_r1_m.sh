#!/bin/bash
# pass M: A-only evict_last hint and intermediate group heights, DRAM traffic only.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; }
for cfg in 1:16 1:24 1:32 0:24 0:12; do
  export VC_GEMM_L2_HINT=${cfg%%:*} VC_GEMM_GROUP_M=${cfg##*:}
  T=100 run m_traffic_h${VC_GEMM_L2_HINT}_g${VC_GEMM_GROUP_M} ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_bf16_kernel --csv --log-file gpurun_out/m_traffic_h${VC_GEMM_L2_HINT}_g${VC_GEMM_GROUP_M}.csv python tools/gemm_traffic.py
done
