#!/bin/bash
# round 2, 2-GPU diagnostic call M: where does the fused sequence-parallel forward leave the single-GPU forward on the 4-layer sp_check model?
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for cfg in "4 9" "4 7" "2 9" "6 9"; do
  set -- $cfg
  SPD_LAYERS=$1 VC_XATTN_VARIANT=$2 timeout 200 $TR --master-port 2961$1 tools/sp_diag.py 2>&1 | grep "^\[" | sed "s/^/x$2 /"
done
SPD_LAYERS=4 SPD_DIM=2048 SPD_HEADS=16 timeout 200 $TR --master-port 29620 tools/sp_diag.py 2>&1 | grep "^\[" | sed "s/^/dim2048 /"
