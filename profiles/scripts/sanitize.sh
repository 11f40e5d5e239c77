#!/bin/bash
# compute-sanitizer over the tcgen05 / TMA / mbarrier / PDL kernels at small shapes (VERDICT r01
# item 8).  ONE tool per gpurun call (B200_PROFILING.md): usage  bash profiles/scripts/sanitize.sh memcheck|racecheck
# The log goes to gpurun_out/sanitize_<tool>.log; a summary is kept under profiles/.
tool=${1:-memcheck}
cd "${GRAFT_REPO_ROOT:-.}" || exit 1
mkdir -p gpurun_out
sel='cached_decode_steps or fused_attention_prologue or cross_attention_block or layernorm_fast or skinny_gemm or flash_decode or dac_conv1d_planes or dac_convtr1d_planes or fp8_attention or fast_path_launches'
# the plain run first: a fault under the sanitizer on this driver can take the GPU down (B200_PROFILING.md)
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_api.py -m gpu -q -x -k "$sel" \
  > gpurun_out/sanitize_plain.log 2>&1 || { tail -20 gpurun_out/sanitize_plain.log; echo "plain run failed: not sanitizing"; exit 1; }
tail -2 gpurun_out/sanitize_plain.log
timeout 2400 compute-sanitizer --tool "$tool" --error-exitcode 7 --launch-timeout 120 \
  python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_api.py -m gpu -q -x -k "$sel" \
  > "gpurun_out/sanitize_${tool}.log" 2>&1
echo "compute-sanitizer --tool $tool rc=$?" | tee -a "gpurun_out/sanitize_${tool}.log"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" "gpurun_out/sanitize_${tool}.log" | tail -5
