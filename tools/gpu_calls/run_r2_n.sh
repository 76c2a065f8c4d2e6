#!/bin/bash
# round 2, GPU call N (1 GPU): end-of-round verification of the default paths the driver runs: smoke(), pytest -m gpu, default
# bench.py line (both arms), then the c5 sampling loop on one GPU
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/n_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/n_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/n_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/n_pytest.log
timeout 900 python bench.py 2> gpurun_out/n_bench.err | grep '^{' > gpurun_out/n_bench_c3.json; echo "bench rc=$?"
python - <<'P'
import json
d=json.load(open('gpurun_out/n_bench_c3.json')); k=d['kernels']
print(round(d['ms_per_step'],1), round(d['e2e']['ms_per_step'],1), d['roofline']['frac'], {n:round(v.get('frac_of_bf16_sustained',v.get('frac_of_hbm',0)),3) for n,v in k.items()}, d['clocks'], d.get('cpu_baseline',{}).get('value'))
P
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 2> gpurun_out/n_ref.err | grep '^{' > gpurun_out/n_bench_ref.json; echo "reference arm rc=$?"; cut -c1-300 gpurun_out/n_bench_ref.json
timeout 600 python tools/c5_loop.py --steps 50 2> gpurun_out/n_c5.err | grep '^{' > gpurun_out/n_c5_loop_n1.json; echo "c5 rc=$?"; cut -c1-600 gpurun_out/n_c5_loop_n1.json
