#!/bin/bash
# round 3, GPU call G (1 GPU): ncu launch list of the round-3 bench command (gpu__time_duration only; per-launch times are
# cold-cache and serialised — only the kernels' SHARES of a step are compared with the CUDA-event numbers of bench.py)
mkdir -p gpurun_out
echo "bench (without ncu) exited 0 in the previous attempt of this call: gpurun_out/r3g_bench.json"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 10000 --csv --log-file gpurun_out/r3g_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-gpu-reference > gpurun_out/r3g_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_shares.py gpurun_out/r3g_launches.csv gpurun_out/launches_r03_bench_c3.csv gpurun_out/launch_shares_r03.json; echo "shares rc=$?"
cut -c1-1500 gpurun_out/launch_shares_r03.json
