#!/bin/bash
# Round-1 GPU pass A: parity tests, smoke, default bench (own arm + reference arm).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > gpurun_out/smi.log 2>&1
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -6 gpurun_out/$name.log | cut -c1-2500; }
T=900 run t_gpu python -m pytest tests -m gpu -q -x
T=300 run smoke python __graft_entry__.py
T=900 run bench_c3 python bench.py --steps 2 --warmup 3
T=600 run bench_ref python bench.py --impl reference --steps 2 --warmup 1
