#!/bin/bash
# round 2, GPU call D (2 GPUs): Ulysses P-invariance incl. ragged/padded token counts, bench with sp_parity at the real c3
# shape, single vs CTA-pair GEMM schedule under sequence parallelism (same box, alternating)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/sp_check.py > gpurun_out/sp_check_n2.log 2>&1; echo "sp_check rc=$?"; grep sp_check gpurun_out/sp_check_n2.log
for i in 1 2; do
  for s in single pair; do
    VC_GEMM_SCHED=$s timeout 600 $TR --master-port 2952$i bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/n2_${s}_$i.json 2> gpurun_out/n2_${s}_$i.err; echo "bench n2 $s $i rc=$?"
    python -c "
import json; d=json.load(open('gpurun_out/n2_${s}_$i.json')); print('$s', d['ms_per_step'], d['sp_parity'], round(d['kernels']['gemm']['tflops']), round(d['kernels']['attn_self']['tflops']), d['clocks']['sm_mhz'])" || tail -5 gpurun_out/n2_${s}_$i.err
  done
done
