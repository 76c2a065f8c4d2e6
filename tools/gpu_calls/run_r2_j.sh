#!/bin/bash
# round 2, GPU call J (1 GPU): persistent short-key attention (variant 9) parity + microbench + c3 step A/B; ncu captures of the
# round-2 kernels (patch embedding, CTA-pair GEMM, ring LayerNorm, x3)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_guard_bands_gpu.py -x -q -p no:cacheprovider -k "x2 or x3 or rescale or flash_attn_guard" > gpurun_out/j_pytest_x3.log 2>&1; rc=$?; echo "pytest x3 rc=$rc"; tail -3 gpurun_out/j_pytest_x3.log
if [ $rc -ne 0 ]; then tail -30 gpurun_out/j_pytest_x3.log; exit 1; fi
rm -f gpurun_out/microbench.jsonl
timeout 300 python tools/microbench.py xattn > gpurun_out/j_mb_xattn.log 2>&1; echo "microbench rc=$?"; cut -c1-170 gpurun_out/j_mb_xattn.log
for v in 7 9 7 9; do
  VC_XATTN_VARIANT=$v timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2> gpurun_out/j_bench_x$v.err | grep '^{' > gpurun_out/j_bench_x${v}_$RANDOM.json; echo "bench xattn=$v rc=$?"
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/j_bench_x*.json')):
    d=json.load(open(f)); k=d['kernels']
    print(f, round(d['ms_per_step'],1), round(k['attn_cross']['ms_per_launch'],3), round(k['attn_cross']['tflops']), round(k['attn_self']['tflops']), round(k['gemm']['tflops']), d['clocks']['sm_mhz'], d.get('host_enqueue_ms_per_step'))
P
timeout 300 python tools/microbench.py xattn patch_embed > gpurun_out/j_plain.log 2>&1 && {
  ncu --set full --clock-control none --import-source on -k regex:flash_attn_x3 -c 1 -o gpurun_out/prof_x3_r02 python tools/microbench.py xattn > gpurun_out/j_ncu_x3.log 2>&1; echo "ncu x3 rc=$?"
  ncu --set full --clock-control none --import-source on -k regex:patch_embed_kernel -c 2 -o gpurun_out/prof_patch_embed_r02 python tools/microbench.py patch_embed > gpurun_out/j_ncu_pe.log 2>&1; echo "ncu patch_embed rc=$?"
}
timeout 300 python tools/microbench.py rows_ab > gpurun_out/j_plain2.log 2>&1 && {
  ncu --set full --clock-control none --import-source on -k regex:ln_modulate_ring -c 1 -o gpurun_out/prof_ln_ring_r02 python tools/microbench.py rows_ab > gpurun_out/j_ncu_ln.log 2>&1; echo "ncu ln ring rc=$?"
}
VC_MB_GEMM_ONE=1 timeout 300 python tools/microbench.py gemm_sched > gpurun_out/j_plain3.log 2>&1 && {
  VC_MB_GEMM_ONE=1 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_2cta -c 1 -o gpurun_out/prof_gemm_pair_r02 python tools/microbench.py gemm_sched > gpurun_out/j_ncu_gemm.log 2>&1; echo "ncu gemm pair rc=$?"
}
