#!/bin/bash
# Round-1 pass D: control-map / loop-glue / patch-embed tests, full parity suite with the prefetching row kernels,
# row + patch-embed microbench, c3 and c2 bench lines.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -80 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -8 gpurun_out/$name.log | cut -c1-1800; }
T=900 run t_gpu python -m pytest tests -m gpu -q
T=300 run mb_rows_pe python tools/microbench.py rows patch_embed
T=600 run bench_c3 python bench.py --steps 3 --warmup 3
T=600 run bench_c2 python bench.py --steps 3 --warmup 3 --config c2
ls -la gpurun_out | tail -12
