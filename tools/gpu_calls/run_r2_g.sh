#!/bin/bash
# round 2, GPU call G (2 GPUs): guard-band tests + emulated-rank tests of the bulk-store scatter, then the in-process A/B of
# the exchange schedules at c3 (tools/sp_overlap_ab.py)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_guard_bands_gpu.py tests/test_kernels_gpu.py -x -q -p no:cacheprovider -k "guard or ulysses" > gpurun_out/g_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/g_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29531 tools/sp_overlap_ab.py --steps 2 --warmup 1 2> gpurun_out/g_ab.err | grep '^{' | tee gpurun_out/g_ab.jsonl
tail -3 gpurun_out/g_ab.err
