#!/bin/bash
# round 2, GPU call F (2 GPUs): overlapped Ulysses exchange (q/k scatter under the next projection, v scattered by its GEMM)
# vs the one-scatter path: emulated-rank bit-identity, P-invariance on 2 real GPUs, same-box ABAB of the c3 step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -p no:cacheprovider -k "ulysses" > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/f_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for ov in 1 0; do
  VC_SP_OVERLAP=$ov timeout 600 $TR --master-port 2951$ov tools/sp_check.py > gpurun_out/f_sp_check_ov$ov.log 2>&1; echo "sp_check ov=$ov rc=$?"; grep "^sp_check" gpurun_out/f_sp_check_ov$ov.log
done
for i in 1 2; do
  for ov in 0 1; do
    VC_SP_OVERLAP=$ov timeout 600 $TR --master-port 2952$i bench.py --gpus 2 --steps 3 --warmup 3 2> gpurun_out/f_n2_ov${ov}_$i.err | grep '^{' > gpurun_out/f_n2_ov${ov}_$i.json; echo "bench n2 ov=$ov $i rc=$?"
    python -c "
import json; d=json.load(open('gpurun_out/f_n2_ov${ov}_$i.json')); k=d['kernels']; print('ov$ov', d['ms_per_step'], d['sp_parity']['cosine'], d['sp_parity']['bit_identical'], {n:round(v['ms_total']/d['steps'],1) for n,v in k.items()}, d['clocks']['sm_mhz'])" || tail -5 gpurun_out/f_n2_ov${ov}_$i.err
  done
done
