#!/bin/bash
# round 2, GPU call T (1 GPU): patch-embedding epilogue software pipeline: tests, the kernel alone, ncu capture
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_guard_bands_gpu.py tests/test_fullsize_gpu.py tests/test_model_gpu.py -x -q -p no:cacheprovider -k "patch or embed or tiny or c1 or golden or guard" > gpurun_out/t_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -3 gpurun_out/t_pytest.log
if [ $rc -ne 0 ]; then tail -40 gpurun_out/t_pytest.log; exit 1; fi
rm -f gpurun_out/microbench.jsonl
timeout 300 python tools/microbench.py patch_embed > gpurun_out/t_pe.log 2>&1; echo "microbench rc=$?"; cut -c1-260 gpurun_out/t_pe.log && {
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:patch_embed_kernel -c 2 -o gpurun_out/prof_patch_embed_r02t python tools/microbench.py patch_embed > gpurun_out/t_ncu_pe.log 2>&1; echo "ncu rc=$?"
}
