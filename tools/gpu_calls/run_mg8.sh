#!/bin/bash
# 8-GPU pass: P-invariance at world 2/4/8 (small model, both exchange modes), then the c3 headline bench and the c4
# long-sequence stress at 8 GPUs (fused peer-memory exchange).
N=${1:-8}
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -60 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -6 gpurun_out/$name.log | cut -c1-3000; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
nvidia-smi topo -m > gpurun_out/topo_$N.log 2>&1
T=400 run t_sp python -m pytest tests/test_sp_gpu.py -m gpu -q
T=300 run bench_c3_n${N} $TR bench.py --gpus $N --steps 3 --warmup 3
T=300 run bench_c4_n${N} $TR bench.py --gpus $N --steps 2 --warmup 2 --config c4
