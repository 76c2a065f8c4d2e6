#!/bin/bash
# round 2, GPU call O (1 GPU): packed-fp32x2 LayerNorm row body: kernel / guard-band / model parity tests, the row kernels alone,
# one c3 bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_guard_bands_gpu.py tests/test_model_gpu.py tests/test_fullsize_gpu.py tests/test_reference_gpu.py -x -q -p no:cacheprovider > gpurun_out/o_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -3 gpurun_out/o_pytest.log
if [ $rc -ne 0 ]; then tail -40 gpurun_out/o_pytest.log; exit 1; fi
rm -f gpurun_out/microbench.jsonl
timeout 300 python tools/microbench.py rows_ab > gpurun_out/o_rows.log 2>&1; echo "rows rc=$?"; grep "ln_" gpurun_out/o_rows.log | cut -c1-330
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2> gpurun_out/o_bench.err | grep '^{' > gpurun_out/o_bench_c3.json; echo "bench rc=$?"
python - <<'P'
import json
d=json.load(open('gpurun_out/o_bench_c3.json')); k=d['kernels']
print(round(d['ms_per_step'],1), {n:(round(v['ms_per_launch'],4), round(v.get('frac_of_bf16_sustained',v.get('frac_of_hbm',0)),3)) for n,v in k.items()}, d['clocks'])
P
