#!/bin/bash
# round 2, GPU call U (1 GPU): x3 epilogue with pipelined TMEM loads: attention tests, the variants alone, one c3 bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_guard_bands_gpu.py -x -q -p no:cacheprovider -k "flash or attn or x3 or x2" > gpurun_out/u_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -2 gpurun_out/u_pytest.log
if [ $rc -ne 0 ]; then tail -40 gpurun_out/u_pytest.log; exit 1; fi
rm -f gpurun_out/microbench.jsonl
timeout 300 python tools/microbench.py xattn > gpurun_out/u_mb_xattn.log 2>&1; echo "microbench rc=$?"; grep "v9\|v7" gpurun_out/u_mb_xattn.log | cut -c1-170
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2> gpurun_out/u_bench.err | grep '^{' > gpurun_out/u_bench_c3.json; echo "bench rc=$?"
python - <<'P'
import json
d=json.load(open('gpurun_out/u_bench_c3.json')); k=d['kernels']
print(round(d['ms_per_step'],1), {n:(round(v['ms_per_launch'],4), round(v.get('frac_of_bf16_sustained',v.get('frac_of_hbm',0)),3)) for n,v in k.items()}, d['clocks'])
P
