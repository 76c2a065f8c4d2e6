#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -12 gpurun_out/$name.log | cut -c1-420; }
VC_ATTN_VARIANTS=1,2,3,4,5,1,2,3,4,5 T=300 run mb_attn_sus python tools/microbench.py attn_sustained
