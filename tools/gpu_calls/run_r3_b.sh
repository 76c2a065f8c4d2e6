#!/bin/bash
# round 3, GPU call B (1 GPU): reference-itself tests (padded keys, c2 full size), c3 full-size reference run, default bench
# line with gpu_reference + the reference arm with the step-independent CPU sample
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_reference_gpu.py tests/test_sp_gpu.py tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r3b_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r3b_pytest.log
cat gpurun_out/ref_gpu_full_c2_test.json | cut -c1-1200
timeout 900 python tools/ref_gpu_full.py --config c3 --iters 3 > gpurun_out/r3b_ref_c3.log 2>&1; echo "ref c3 rc=$?"; tail -3 gpurun_out/r3b_ref_c3.log | cut -c1-1500
timeout 900 python bench.py --steps 5 --warmup 3 2> gpurun_out/r3b_bench.err | grep '^{' > gpurun_out/r3b_bench.json; echo "bench rc=$?"
python - <<P
import json
d=json.load(open('gpurun_out/r3b_bench.json'))
print(round(d['ms_per_step'],1), round(d['e2e']['ms_per_step'],1), d.get('gpu_reference'), d.get('cpu_baseline',{}).get('value'), d['clocks'])
P
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 2> gpurun_out/r3b_refarm.err | grep '^{' > gpurun_out/r3b_refarm.json; echo "refarm rc=$?"; cut -c1-300 gpurun_out/r3b_refarm.json
