#!/bin/bash
# round 2, GPU call C: ring row kernels + 8-warp GEMM epilogue: parity, A/B microbenchmarks, c3 and c2 step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_reference_gpu.py tests/test_fullsize_gpu.py tests/test_model_gpu.py -x -q > gpurun_out/pytest_c.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_c.log
cat gpurun_out/reference_gpu_parity.jsonl | cut -c1-900
timeout 300 python tools/microbench.py rows_ab > gpurun_out/rows_ab.log 2>&1; echo "rows_ab rc=$?"; cut -c1-500 gpurun_out/rows_ab.log | tail -8
timeout 600 python tools/microbench.py gemm_sched > gpurun_out/gemm_sched_c.log 2>&1; echo "gemm_sched rc=$?"; python - <<'PY'
import json
for ln in open('gpurun_out/gemm_sched_c.log'):
    if ln.startswith('{'):
        d=json.loads(ln); print(d['shape'], d['kernel'], {k: round(d[k]['tflops']) for k in ('single','mc','pair','cublaslt')}, d['bit_identical'])
PY
for cfg in c3 c2; do
  timeout 400 python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${cfg}_c.json 2> gpurun_out/bench_${cfg}_c.err; echo "bench $cfg rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/bench_${cfg}_c.json')); print('$cfg', d['ms_per_step'], d['clocks']['sm_mhz'], {k:(round(v.get('tflops',v.get('gbs',0))), round(v['ms_per_launch'],3)) for k,v in d['kernels'].items()})"
done
VC_GEMM_SCHED=pair timeout 400 python bench.py --config c2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_c_pair.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_c2_c_pair.json')); print('c2 pair', d['ms_per_step'], d['clocks']['sm_mhz'], round(d['kernels']['gemm']['tflops']))"
VC_ROW_KERNELS=regs timeout 400 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_c_regs.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bench_c3_c_regs.json')); print('c3 regs', d['ms_per_step'], d['clocks']['sm_mhz'], {k:(round(v.get('tflops',v.get('gbs',0))), round(v['ms_per_launch'],3)) for k,v in d['kernels'].items()})"
