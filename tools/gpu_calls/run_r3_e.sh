#!/bin/bash
# round 3, GPU call E (8 GPUs): final-tree multi-GPU records — P-invariance tests at world 2/4/8 (barrier time-out changes),
# c3 at N=8, the c5 50-step two-expert loop at N=8, c4 at N=8/4/2.  Ordered by priority; every command has its own limit.
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; ( timeout -k 10 $T "$@" 2>gpurun_out/$name.err | tail -40 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -3 gpurun_out/$name.log | cut -c1-1500; }
tr() { echo "python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2"; }
T=420 run r3e_sp python -m pytest tests/test_sp_gpu.py -m gpu -q -p no:cacheprovider
T=240 run r3e_bench_c3_n8 $(tr 8 29611) bench.py --gpus 8 --steps 5 --warmup 3
T=300 run r3e_c5_loop_n8 $(tr 8 29612) tools/c5_loop.py --steps 50
T=240 run r3e_bench_c4_n8 $(tr 8 29613) bench.py --gpus 8 --steps 2 --warmup 3 --config c4
T=240 run r3e_bench_c4_n4 $(tr 4 29614) bench.py --gpus 4 --steps 2 --warmup 3 --config c4
T=300 run r3e_bench_c4_n2 $(tr 2 29615) bench.py --gpus 2 --steps 2 --warmup 3 --config c4
