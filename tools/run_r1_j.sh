#!/bin/bash
# pass J (1 GPU): ncu evidence for the round-1 end state — launch list of one bench run (all kernels of this repo), then
# full captures of the row kernels, the GEMM, and the short-key attention kernel at their c3 launch shapes.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -30 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -3 gpurun_out/$name.log | cut -c1-600; }
K='regex:flash_attn|gemm_bf16_kernel|ln_modulate|rmsnorm_rope|patchify_kernel|patch_embed_kernel'
T=700 run j_launches ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file gpurun_out/launches_r01j.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline
T=200 run j_ncu_ln ncu --set full --clock-control none --import-source on -k regex:ln_modulate_block_kernel -s 3 -c 1 -f -o gpurun_out/prof_ln_r01j python tools/microbench.py rows
T=200 run j_ncu_rms ncu --set full --clock-control none --import-source on -k regex:rmsnorm_rope_block_kernel -s 3 -c 1 -f -o gpurun_out/prof_rms_r01j python tools/microbench.py rows
T=200 run j_ncu_x1 ncu --set full --clock-control none --import-source on -k regex:flash_attn_x1_kernel -s 1 -c 1 -f -o gpurun_out/prof_x1_r01j python tools/microbench.py xattn
T=300 run j_ncu_gemm ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 1 -c 1 -f -o gpurun_out/prof_gemm_r01j python tools/microbench.py gemm
T=200 run j_mb_rows python tools/microbench.py rows gemm
ls -la gpurun_out | tail -12
