#!/bin/bash
# 8-GPU pass B: P-invariance at world 8 (both exchange modes), c3 bench at N=2/4/8, c5 50-step loop at N=8.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -40 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -3 gpurun_out/$name.log | cut -c1-2500; }
TR() { echo python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2; }
T=240 run mg_sp_check_8 $(TR 8 29611) tools/sp_check.py
T=400 run mg_c5_n8 $(TR 8 29612) tools/c5_loop.py --steps 50
T=300 run mg_bench_c3_n8 $(TR 8 29613) bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline
T=300 run mg_bench_c3_n4 $(TR 4 29614) bench.py --gpus 4 --steps 3 --warmup 3 --no-cpu-baseline
T=300 run mg_bench_c3_n2 $(TR 2 29615) bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline
