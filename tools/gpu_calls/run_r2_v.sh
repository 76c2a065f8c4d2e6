#!/bin/bash
# round 2, GPU call V (1 GPU): final tree: smoke, full GPU suite, c2 and c3 bench lines
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/v_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/v_pytest.log
for c in c2 c3; do
  timeout 900 python bench.py --config $c --steps 2 --warmup 3 --no-cpu-baseline 2> gpurun_out/v_bench_$c.err | grep '^{' > gpurun_out/v_bench_$c.json; echo "bench $c rc=$?"
  python - <<P
import json
d=json.load(open('gpurun_out/v_bench_$c.json')); k=d['kernels']
print('$c', round(d['ms_per_step'],1), round(d['e2e']['ms_per_step'],1), round(d['step_frac_of_bf16_sustained'],3), {n:(round(v['ms_per_launch'],4), round(v.get('frac_of_bf16_sustained',v.get('frac_of_hbm',0)),3)) for n,v in k.items()}, d['clocks'])
P
done
