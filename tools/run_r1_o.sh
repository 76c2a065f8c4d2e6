#!/bin/bash
# pass O: end-of-round verification on one B200 — parity suite, smoke, default bench (both arms).
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -2 gpurun_out/$name.log | cut -c1-400; }
T=600 run o_tests python -m pytest tests -m gpu -q -x
T=200 run o_smoke python __graft_entry__.py
T=400 run o_bench python bench.py
T=400 run o_bench_ref python bench.py --impl reference
