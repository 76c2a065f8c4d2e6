#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -8 gpurun_out/$name.log | cut -c1-1800; }
T=600 run t_rows python -m pytest tests/test_kernels_gpu.py -q -x -k "ln_ or rmsnorm"
T=300 run mb_rows python tools/microbench.py rows
