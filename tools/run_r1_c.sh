#!/bin/bash
# Round-1 pass C: new kernels (implicit-GEMM patch embedding, loop glue, expert switch) + full parity suite, then ncu
# evidence for the default bench command and the dominant kernel at its c3 launch shape.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -8 gpurun_out/$name.log | cut -c1-1800; }
T=200 run t_new python -m pytest tests/test_kernels_gpu.py tests/test_loop_glue.py tests/test_expert_switch.py -m gpu -q -x -k "patch_embed or loop or control or cfg_euler or denoise or expert or routed"
T=900 run t_gpu python -m pytest tests -m gpu -q
T=300 run smoke python __graft_entry__.py
T=600 run bench_c3 python bench.py --steps 2 --warmup 3
K='regex:flash_attn|gemm_bf16_kernel|ln_modulate|rmsnorm_rope|patchify_kernel|patch_embed_kernel'
T=900 run ncu_launches ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 2793 -c 1862 --csv --log-file gpurun_out/launches_r01c.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline
VC_ATTN_VARIANTS=4 T=600 run ncu_attn ncu --set full --clock-control none --import-source on -k regex:flash_attn_sa_kernel -s 1 -c 1 -f -o gpurun_out/prof_attn_c3_r01 \
    python tools/microbench.py attn_c3
T=300 run ncu_pe ncu --set full --clock-control none --import-source on -k regex:patch_embed_kernel -s 1 -c 1 -f -o gpurun_out/prof_patch_embed_r01 \
    python tools/microbench.py patch_embed
T=200 run mb_pe python tools/microbench.py patch_embed
ls -la gpurun_out | tail -24
