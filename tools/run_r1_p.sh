#!/bin/bash
# pass P: cluster-multicast GEMM variant (VC_GEMM_MC=1): parity tests, then microbench with and without.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 5 $T "$@" 2>&1 | tail -25 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -4 gpurun_out/$name.log | cut -c1-700; }
VC_GEMM_MC=1 T=60 run p_tests_mc python -m pytest tests/test_kernels_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x -k gemm
if grep -q " passed" gpurun_out/p_tests_mc.log && ! grep -q "failed\|error" gpurun_out/p_tests_mc.log; then
  VC_GEMM_MC=1 T=60 run p_mb_mc python tools/microbench.py gemm
  T=60 run p_mb_plain python tools/microbench.py gemm
fi
