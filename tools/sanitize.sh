#!/bin/bash
# compute-sanitizer over smoke-sized runs of every hand-rolled mbarrier / cluster protocol (GPU box only):
#   memcheck  : out-of-bounds / misaligned global + shared accesses, leaked TMEM, invalid TMA boxes
#   racecheck : shared-memory hazards between the warp roles (generic-proxy accesses; async-proxy traffic — TMA,
#               tcgen05 — is outside racecheck's view and is ordered by the mbarriers themselves)
# Usage (from the repo root):  gpurun -- 'bash tools/sanitize.sh > gpurun_out/sanitize.log 2>&1'
# The test selection keeps sizes small: the sanitizer runs kernels 10-100x slower.
set -u
CS=${CS:-/usr/local/cuda/bin/compute-sanitizer}
PY=${PY:-python}
SEL='test_gemm_tn or test_gemm_epilogues or (test_gemm_fused_geglu_fwd_bwd and not 22496) or (test_attn_gemma_fwd and (200 or 40-)) or (test_attn_gemma_bwd and (200 or 40-)) or (test_attn_siglip_fwd and (100 or 9-))'
FILES="tests/test_gemm_gpu.py tests/test_attn_gpu.py"
rc=0
for tool in memcheck racecheck; do
  echo "===== compute-sanitizer --tool $tool"
  timeout -k 10 1500 $CS --tool $tool --error-exitcode 3 --print-limit 20 \
      $PY -m pytest $FILES -q -x -m gpu -k "$SEL" -p no:cacheprovider 2>&1 | tail -40
  r=${PIPESTATUS[0]}
  echo "===== $tool exit code $r"
  [ "$r" -ne 0 ] && rc=$r
done
echo "===== norm / misc kernels under memcheck"
timeout -k 10 900 $CS --tool memcheck --error-exitcode 3 --print-limit 20 \
    $PY -m pytest tests/test_norm_gpu.py tests/test_misc_gpu.py -q -x -m gpu -k "575 or 703 or 64 or splice or colsum" -p no:cacheprovider 2>&1 | tail -15
r=${PIPESTATUS[0]}; echo "===== memcheck (bandwidth kernels) exit code $r"; [ "$r" -ne 0 ] && rc=$r
exit $rc
