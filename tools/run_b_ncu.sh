#!/bin/bash
# Round-1 GPU pass B: ncu evidence. (1) launch list of the default bench command (timed region only),
# (2) one full capture per kernel class on the microbench commands (c3 shapes).
mkdir -p gpurun_out
K='regex:flash_attn_kernel|gemm_bf16_kernel|ln_modulate_kernel|rmsnorm_rope_kernel|patchify_kernel'
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 2805 -c 1870 --csv --log-file gpurun_out/launches_r01.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ncu.log 2>&1
echo "launch list rc=$?"
VC_ATTN_VARIANTS=${VC_ATTN_VARIANT:-1} python tools/microbench.py attn > gpurun_out/mb_attn_plain.log 2>&1 &&
VC_ATTN_VARIANTS=${VC_ATTN_VARIANT:-1} ncu --set full --clock-control none --import-source on -k regex:flash_attn_kernel -s 2 -c 1 -f -o gpurun_out/prof_attn_r01 \
    python tools/microbench.py attn > gpurun_out/mb_attn_ncu.log 2>&1
echo "attn full rc=$?"
python tools/microbench.py rows > gpurun_out/mb_rows_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ln_modulate_kernel -s 3 -c 1 -f -o gpurun_out/prof_ln_r01 \
    python tools/microbench.py rows > gpurun_out/mb_rows_ncu.log 2>&1
echo "ln full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:rmsnorm_rope_kernel -s 3 -c 1 -f -o gpurun_out/prof_rms_r01 \
    python tools/microbench.py rows > gpurun_out/mb_rows_ncu2.log 2>&1
echo "rms full rc=$?"
python tools/microbench.py gemm > gpurun_out/mb_gemm_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 3 -c 1 -f -o gpurun_out/prof_gemm_r01 \
    python tools/microbench.py gemm > gpurun_out/mb_gemm_ncu.log 2>&1
echo "gemm full rc=$?"
ls -la gpurun_out | tail -20
