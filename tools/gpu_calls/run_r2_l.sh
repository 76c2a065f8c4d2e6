#!/bin/bash
# round 2, multi-GPU call L (N = $1 GPUs, default 2): P-invariance incl. the interleaved two-chain schedule, then an in-process
# ABAB of the plain schedule vs the interleaved one at c3 (model built once), then one bench line with the faster one
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_guard_bands_gpu.py -x -q -p no:cacheprovider > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/l_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29601 tools/sp_check.py > gpurun_out/l${N}_sp_check.log 2>&1; rc=$?; echo "sp_check rc=$rc"; grep "^sp_check" gpurun_out/l${N}_sp_check.log
if [ $rc -ne 0 ]; then tail -30 gpurun_out/l${N}_sp_check.log; exit 1; fi
timeout 900 $TR --master-port 29602 tools/sp_overlap_ab.py --config ${2:-c3} --steps ${3:-3} --warmup 2 --seq one_scatter,interleave,one_scatter,interleave,one_scatter,interleave > gpurun_out/l${N}_ab.log 2> gpurun_out/l${N}_ab.err; echo "ab rc=$?"
grep '^{' gpurun_out/l${N}_ab.log > gpurun_out/sp_interleave_ab_n${N}.jsonl; cut -c1-400 gpurun_out/sp_interleave_ab_n${N}.jsonl; tail -5 gpurun_out/l${N}_ab.err
