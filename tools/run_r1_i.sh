#!/bin/bash
# pass I (1 GPU): short-key attention kernel (variant 6) parity + microbench vs variant 4; eager PyTorch+FA2 block bar.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -40 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -6 gpurun_out/$name.log | cut -c1-1500; }
T=120 run i_tests python -m pytest tests/test_kernels_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x -k "short_keys or rescale or cross_attention"
T=90 run i_xattn python tools/microbench.py xattn
T=150 run i_refblock python tools/ref_gpu_block.py
