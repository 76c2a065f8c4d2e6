#!/bin/bash
# Round-1 pass G on one B200: full GPU parity suite (UniPC + step-cache tests are new), smoke, default bench.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -60 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -5 gpurun_out/$name.log | cut -c1-3000; }
T=1200 run g_tests python -m pytest tests -m gpu -q -x --durations=8
T=300 run g_smoke python __graft_entry__.py
T=900 run g_bench_c3 python bench.py --steps 3 --warmup 3
