#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -12 gpurun_out/$name.log | cut -c1-600; }
T=300 run t_fused python -m pytest tests/test_kernels_gpu.py -q -x -k "ulysses or rmsnorm"
