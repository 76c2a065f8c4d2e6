#!/bin/bash
# round 2, final multi-GPU call Q (N = $1 GPUs, default 8): P-invariance in three modes, then the c3 bench line of the end-of-round tree
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29601 tools/sp_check.py > gpurun_out/q${N}_sp_check.log 2>&1; echo "sp_check rc=$?"; grep "^sp_check" gpurun_out/q${N}_sp_check.log
timeout 600 $TR --master-port 29602 bench.py --gpus $N --config c3 --steps 3 --warmup 3 2> gpurun_out/q${N}_bench_c3.err | grep '^{' > gpurun_out/q${N}_bench_c3.json; echo "bench c3 rc=$?"
python - <<P
import json
d=json.load(open('gpurun_out/q${N}_bench_c3.json')); k=d['kernels']
print('c3 N=$N', round(d['ms_per_step'],1), round(d['e2e']['ms_per_step'],1), d['sp_parity'], {n:round(v['ms_total']/d['steps'],1) for n,v in k.items()}, d['clocks'])
P
