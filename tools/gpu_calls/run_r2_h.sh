#!/bin/bash
# round 2, GPU call H (2 GPUs): own epoch-flag barrier kernel vs torch symmetric-memory barrier; P-invariance with the own barrier
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for b in own torch; do
  VC_SP_BARRIER=$b timeout 300 $TR --master-port 2954${#b} tools/sp_check.py > gpurun_out/h_sp_check_$b.log 2>&1; echo "sp_check barrier=$b rc=$?"; grep "^sp_check" gpurun_out/h_sp_check_$b.log | tail -2
done
timeout 900 $TR --master-port 29533 tools/sp_overlap_ab.py --steps 2 --warmup 1 --only one_scatter_torch_barrier,one_scatter,full 2> gpurun_out/h_ab.err | grep '^{' | tee gpurun_out/h_ab.jsonl
tail -3 gpurun_out/h_ab.err
