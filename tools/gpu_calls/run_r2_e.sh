#!/bin/bash
# round 2, GPU call E (1 GPU): full GPU suite (incl. the reference-pipeline boundary tests), compute-sanitizer pass, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/e_pytest.log
SAN_TIMEOUT=500 bash tools/sanitize.sh
timeout 600 python bench.py > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "bench rc=$?"; head -c 600 gpurun_out/e_bench.json
