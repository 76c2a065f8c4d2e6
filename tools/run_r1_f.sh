#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -60 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -12 gpurun_out/$name.log | cut -c1-1500; }
T=300 run t_raster python -m pytest tests/test_raster.py tests/test_control_maps.py -m gpu -q
T=900 run t_gpu python -m pytest tests -m gpu -q
T=300 run mb_rows python tools/microbench.py rows
T=600 run bench_raster python tools/bench_raster.py
