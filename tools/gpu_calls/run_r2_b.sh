#!/bin/bash
# round 2, GPU call B: CTA-pair (cta_group::2) GEMM parity + schedule microbench + same-box ABAB of the c3 step,
# after the full GPU suite (now with the c2 and c3 full-depth parity tests and the a4 kernels)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "pair or time_embed or modulation or ulysses or rejects or multicast" > gpurun_out/pytest_b_pair.log 2>&1; echo "pair tests rc=$?"; tail -3 gpurun_out/pytest_b_pair.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_b.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_b.log
timeout 900 python tools/microbench.py gemm_sched > gpurun_out/gemm_sched.log 2>&1; echo "gemm_sched rc=$?"; tail -12 gpurun_out/gemm_sched.log | cut -c1-700
for i in 1 2; do
  for s in single pair; do
    VC_GEMM_SCHED=$s timeout 400 python bench.py --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/abab_${s}_$i.json 2> gpurun_out/abab_${s}_$i.err; echo "bench $s $i rc=$?"
    python -c "
import json; d=json.load(open('gpurun_out/abab_${s}_$i.json')); print('$s', d['ms_per_step'], d['kernels']['gemm']['tflops'], d['kernels']['attn_self']['tflops'], d['clocks']['sm_mhz'])"
  done
done
