#!/bin/bash
# round 3, GPU call C (1 GPU): TeaCache / cfg-skip pipeline parity, c4 full-depth oracle parity, the reference model at c3
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline_boundary.py tests/test_model_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r3c_pipeline.log 2>&1; echo "pipeline rc=$?"; tail -12 gpurun_out/r3c_pipeline.log | cut -c1-400
timeout 900 python -m pytest tests/test_parity_full_gpu.py -m gpu -x -q -p no:cacheprovider -k c4 > gpurun_out/r3c_c4.log 2>&1; echo "c4 rc=$?"; tail -8 gpurun_out/r3c_c4.log | cut -c1-600
cut -c1-1500 gpurun_out/parity_c4_test.json | head -40
timeout 1200 python tools/ref_gpu_full.py --config c3 --iters 3 > gpurun_out/r3c_ref_c3.log 2>&1; echo "ref c3 rc=$?"; tail -3 gpurun_out/r3c_ref_c3.log | cut -c1-1800
