#!/bin/bash
# round 2, GPU call A: full GPU suite (incl. full-depth c2 parity and reference-on-GPU pins), c3 full-depth parity with the
# fp32 oracle, the reference-module bar, and a short c3 bench line of this box before any kernel change
mkdir -p gpurun_out
export VC_SKIP_C3_PARITY=1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_a.log
tail -5 gpurun_out/pytest_a.log
timeout 900 python tools/parity_full.py --config c3 --fp32 --out gpurun_out/parity_c3.json > gpurun_out/parity_c3.log 2>&1; echo "parity c3 rc=$?"
tail -8 gpurun_out/parity_c3.log
timeout 300 python tools/ref_gpu_block.py > gpurun_out/ref_gpu_block.log 2>&1; echo "ref block rc=$?"; tail -2 gpurun_out/ref_gpu_block.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_a.json 2> gpurun_out/bench_c3_a.err; echo "bench rc=$?"
head -c 600 gpurun_out/bench_c3_a.json
