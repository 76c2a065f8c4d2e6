#!/bin/bash
# round 2, GPU call W (1 GPU): ncu --set full capture with source of the self-attention kernel at the c3 launch shape (round-2 evidence)
mkdir -p gpurun_out
rm -f gpurun_out/microbench.jsonl
timeout 300 python tools/microbench.py attn_c3 > gpurun_out/w_plain.log 2>&1; echo "plain rc=$?"; cut -c1-200 gpurun_out/w_plain.log && {
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:flash_attn_sa -c 1 -o gpurun_out/prof_attn_sa_r02 python tools/microbench.py attn_c3 > gpurun_out/w_ncu.log 2>&1; echo "ncu rc=$?"
}
