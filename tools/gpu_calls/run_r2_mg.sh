#!/bin/bash
# round 2, multi-GPU call (N = $1 GPUs): P-invariance check, bench lines of c3 and c4 with sp_parity, barrier / GEMM-schedule A/B
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29601 tools/sp_check.py > gpurun_out/mg${N}_sp_check.log 2>&1; echo "sp_check rc=$?"; grep "^sp_check" gpurun_out/mg${N}_sp_check.log
for c in c3 c4; do
  timeout 900 $TR --master-port 29602 bench.py --gpus $N --config $c --steps 3 --warmup 3 2> gpurun_out/mg${N}_bench_$c.err | grep '^{' > gpurun_out/mg${N}_bench_$c.json; echo "bench $c rc=$?"
  python - <<P
import json
d=json.load(open('gpurun_out/mg${N}_bench_$c.json')); k=d['kernels']
print('$c N=$N', round(d['ms_per_step'],1), d['sp_parity'], {n:round(v['ms_total']/d['steps'],1) for n,v in k.items()}, d['clocks']['sm_mhz'])
P
done
if [ "$2" == "ab" ]; then
  VC_GEMM_SCHED=single timeout 600 $TR --master-port 29603 bench.py --gpus $N --config c3 --steps 4 --warmup 3 2> gpurun_out/mg${N}_bench_c3_single.err | grep '^{' > gpurun_out/mg${N}_bench_c3_single.json; echo "bench c3 single rc=$?"
  VC_SP_BARRIER=torch timeout 600 $TR --master-port 29604 bench.py --gpus $N --config c3 --steps 4 --warmup 3 2> gpurun_out/mg${N}_bench_c3_torchbar.err | grep '^{' > gpurun_out/mg${N}_bench_c3_torchbar.json; echo "bench c3 torch-barrier rc=$?"
  python - <<P
import json
for n in ('single','torchbar'):
    d=json.load(open('gpurun_out/mg${N}_bench_c3_%s.json'%n)); k=d['kernels']
    print(n, round(d['ms_per_step'],1), {m:round(v['ms_total']/d['steps'],1) for m,v in k.items()}, d['clocks']['sm_mhz'])
P
fi
