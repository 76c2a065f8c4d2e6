#!/bin/bash
# compute-sanitizer pass over one small case of every kernel family (SURVEY §5 "race detection"; VERDICT r1 "next" #7).
# memcheck: out-of-bounds / misaligned global + shared accesses; racecheck: shared-memory hazards between the warp roles
# (hand-rolled mbarrier rings); synccheck: barrier misuse.  Logs -> gpurun_out/sanitizer_<tool>.log.
# CLOSED on this GPU pool (profiles/sanitizer_r02.txt: rc=86, and runs under it left GPUs needing a reset).  It only runs with an
# explicit opt-in on a machine you own; tests/test_guard_bands_gpu.py is the stand-in that runs everywhere.
if [ "${VC_ALLOW_SANITIZER:-0}" != "1" ]; then
  echo "tools/sanitize.sh: compute-sanitizer is closed on the shared pool; set VC_ALLOW_SANITIZER=1 on your own machine" >&2
  exit 2
fi
mkdir -p gpurun_out
SEL='test_gemm_bias and (128-256-64 or 300-512-512 or 77-64-128) or test_gemm_epilogues or (test_gemm_cta_pair_schedule and (650-768-512 or 129-256-64)) or (test_gemm_multicast and 650-768-512) or test_gemm_inplace or (test_flash_attn_shared_s and 333-333 and (4-False or 2-True)) or (test_flash_attn_short_keys_x1 and 333-512) or (test_flash_attn_tmem_p and 256-256) or (test_flash_attn_smem_p and 2-2-256-256) or test_ln_affine or test_ln_modulate_wide_rows or test_rmsnorm_rope_wide_rows or test_rmsnorm_only or (test_row_ring and (7-1536 or 3-5120)) or (test_ulysses and 2-16) or (test_patch_embed_implicit_gemm and 16-2-12-104-256) or test_patchify or (test_time_embed and 2-512) or test_modulation_table'
for tool in memcheck racecheck synccheck; do
  timeout ${SAN_TIMEOUT:-420} compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 7 \
    python -m pytest tests/test_kernels_gpu.py -q -x -p no:cacheprovider -k "$SEL" > gpurun_out/sanitizer_$tool.log 2>&1
  echo "sanitizer $tool rc=$?" | tee -a gpurun_out/sanitizer_$tool.log
  grep -E "ERROR SUMMARY|passed|failed|Error|hazard" gpurun_out/sanitizer_$tool.log | tail -4
done
