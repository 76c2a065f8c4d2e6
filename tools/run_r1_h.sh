#!/bin/bash
# pass H (1 GPU): c5 loop plumbing (small), then 4 full-size steps to check memory with two resident experts.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -40 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -4 gpurun_out/$name.log | cut -c1-3000; }
T=300 run h_c5_small python tools/c5_loop.py --small --steps 8
T=600 run h_c5_full4 python tools/c5_loop.py --steps 4
