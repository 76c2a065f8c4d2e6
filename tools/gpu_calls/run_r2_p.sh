#!/bin/bash
# round 2, multi-GPU call P (N = $1 GPUs, default 2): bench lines of the configs in $2.. (default c4 c3) in the end-of-round state
N=${1:-2}; shift
CFGS=${@:-c4 c3}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for c in $CFGS; do
  timeout 900 $TR --master-port 29602 bench.py --gpus $N --config $c --steps 2 --warmup 3 2> gpurun_out/p${N}_bench_$c.err | grep '^{' > gpurun_out/p${N}_bench_$c.json; echo "bench $c rc=$?"
  python - <<P
import json
d=json.load(open('gpurun_out/p${N}_bench_$c.json')); k=d['kernels']
print('$c N=$N', round(d['ms_per_step'],1), round(d['e2e']['ms_per_step'],1), d['sp_parity'], {n:round(v['ms_total']/d['steps'],1) for n,v in k.items()}, d['clocks'])
P
done
