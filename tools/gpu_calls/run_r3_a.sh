#!/bin/bash
# round 3, GPU call A (1 GPU): the static-scratch time-embed entry point alone in a fresh process, then the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r3a_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/r3a_zz_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q -p no:cacheprovider --durations=15 > gpurun_out/r3a_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r3a_pytest.log
