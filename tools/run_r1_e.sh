#!/bin/bash
# Round-1 pass E: packed-bf16 row kernels — parity (kernel, model, full-size suites), row microbench, c3 bench.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -8 gpurun_out/$name.log | cut -c1-1800; }
T=900 run t_gpu python -m pytest tests -m gpu -q
T=300 run mb_rows python tools/microbench.py rows
T=600 run bench_c3 python bench.py --steps 3 --warmup 3 --no-cpu-baseline
