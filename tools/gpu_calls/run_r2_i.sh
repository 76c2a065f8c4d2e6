#!/bin/bash
# round 2, GPU call I (1 GPU): short-key attention x2 (variants 7/8) parity + microbench + c3 step A/B, demo scene, full suite, ncu of x2
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -p no:cacheprovider -k "x2 or rescale" > gpurun_out/i_pytest_x2.log 2>&1; echo "pytest x2 rc=$?"; tail -3 gpurun_out/i_pytest_x2.log
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/i_pytest.log
rm -f gpurun_out/microbench.jsonl
timeout 300 python tools/microbench.py xattn > gpurun_out/i_mb_xattn.log 2>&1; echo "microbench rc=$?"; cat gpurun_out/i_mb_xattn.log | cut -c1-200
for v in 6 7 8 6 8; do
  VC_XATTN_VARIANT=$v timeout 600 python bench.py --steps 2 --warmup 3 2> gpurun_out/i_bench_x$v.err | grep '^{' > gpurun_out/i_bench_x${v}_$RANDOM.json; echo "bench xattn=$v rc=$?"
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/i_bench_x*.json')):
    d=json.load(open(f)); k=d['kernels']
    print(f, round(d['ms_per_step'],1), round(k['attn_cross']['ms_per_launch'],3), round(k['attn_cross']['tflops']), round(k['attn_self']['tflops']), round(k['gemm']['tflops']), d['clocks']['sm_mhz'])
P
VC_MB_ONLY=1 timeout 300 python tools/microbench.py xattn > gpurun_out/i_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:flash_attn_x2 -c 2 -o gpurun_out/prof_x2_r02 python tools/microbench.py xattn > gpurun_out/i_ncu.log 2>&1; echo "ncu rc=$?"
