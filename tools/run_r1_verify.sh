#!/bin/bash
# Round-1 verification pass on one B200: parity tests, smoke, default bench (both arms), ncu launch list of the timed
# steps, one full ncu capture of the dominant kernel (self-attention at the c3 per-launch shape).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > gpurun_out/smi.log 2>&1
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -6 gpurun_out/$name.log | cut -c1-2500; }
T=900 run t_gpu python -m pytest tests -m gpu -q -x
T=300 run smoke python __graft_entry__.py
T=900 run bench_c3 python bench.py --steps 2 --warmup 3
T=600 run bench_ref python bench.py --impl reference --steps 2 --warmup 1
K='regex:flash_attn_kernel|gemm_bf16_kernel|ln_modulate|rmsnorm_rope|patchify_kernel'
T=900 run ncu_launches ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 2805 -c 1870 --csv --log-file gpurun_out/launches_r01b.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline
VC_ATTN_VARIANTS=4 T=300 run mb_attn python tools/microbench.py attn_c3 attn_sustained
VC_ATTN_VARIANTS=4 T=600 run ncu_attn ncu --set full --clock-control none --import-source on -k regex:flash_attn_kernel -s 1 -c 1 -f -o gpurun_out/prof_attn_c3_r01 \
    python tools/microbench.py attn_c3
ls -la gpurun_out | tail -20
