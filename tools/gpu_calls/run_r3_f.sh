#!/bin/bash
# round 3, GPU call F (1 GPU): final tree — smoke, the whole GPU suite, the default bench line (both arms)
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3f_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r3f_smoke.log | cut -c1-300
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --durations=8 > gpurun_out/r3f_pytest.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/r3f_pytest.log | cut -c1-300
timeout 600 python bench.py --impl reference > gpurun_out/r3f_refarm.json 2> gpurun_out/r3f_refarm.err; echo "refarm rc=$?"; cut -c1-200 gpurun_out/r3f_refarm.json
timeout 900 python bench.py 2> gpurun_out/r3f_bench.err | grep '^{' > gpurun_out/r3f_bench.json; echo "bench rc=$?"
python - <<P
import json
d=json.load(open('gpurun_out/r3f_bench.json')); k=d['kernels']
print(round(d['ms_per_step'],1), round(d['e2e']['ms_per_step'],1), round(d['step_frac_of_bf16_sustained'],3), d['roofline']['frac'], d.get('gpu_reference',{}).get('speedup'), d.get('cpu_baseline',{}).get('value'), d['clocks'])
print({n:(round(v['ms_per_launch'],4), round(v.get('frac_of_bf16_sustained',v.get('frac_of_hbm',0)),3)) for n,v in k.items()})
P
