#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -60 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -12 gpurun_out/$name.log | cut -c1-1500; }
T=600 run t_model python -m pytest tests/test_model_gpu.py -q -x
T=300 run smoke python __graft_entry__.py
T=600 run bench_c1 python bench.py --config c1 --steps 3 --warmup 3
T=900 run bench_c3 python bench.py --steps 2 --warmup 3
