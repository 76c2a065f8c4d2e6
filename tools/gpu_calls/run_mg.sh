#!/bin/bash
# multi-GPU pass: N = $1 GPUs.  P-invariance check (both exchange modes), then the c3 bench in both modes.
N=${1:-2}
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; ( timeout -k 10 $T "$@" 2>&1 | tail -60 ) > gpurun_out/$name.log; echo "exit ${PIPESTATUS[0]}"; tail -6 gpurun_out/$name.log | cut -c1-2500; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
nvidia-smi topo -m > gpurun_out/topo_$N.log 2>&1
T=300 run sp_check_$N $TR tools/sp_check.py
T=600 run bench_c3_n${N}_fused $TR bench.py --gpus $N --steps 2 --warmup 3
VC_SP_MODE=nccl T=600 run bench_c3_n${N}_nccl $TR bench.py --gpus $N --steps 2 --warmup 3
