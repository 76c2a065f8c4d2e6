#!/bin/bash
# First-light run on the GPU box: each group in its own process and timeout so that one hang/fault does not hide the rest.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -40 ) > gpurun_out/$name.log; echo "exit $?"; tail -5 gpurun_out/$name.log; }
T=400 run t_norm python -m pytest tests/test_kernels_gpu.py -q -x -k "ln_ or rmsnorm or patchify or errors"
T=400 run t_gemm python -m pytest tests/test_kernels_gpu.py -q -k "gemm"
T=400 run t_attn0 python -m pytest tests/test_kernels_gpu.py -q -k "smem_p or rescale"
T=400 run t_attn1 python -m pytest tests/test_kernels_gpu.py -q -k "tmem_p"
T=600 run mb_rows_gemm python tools/microbench.py rows gemm
VC_ATTN_VARIANTS=0 T=300 run mb_attn0 python tools/microbench.py attn
VC_ATTN_VARIANTS=1 T=300 run mb_attn1 python tools/microbench.py attn
