#!/bin/bash
# round 2, call 62 (1 GPU): tile GEMM with the K walk of every tile rotated (default) against all tiles in step (norot):
# phases of generate() on one box, GEMM tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_t5.py -m gpu -x -q -k "large_m or t5 or prefill" > gpurun_out/r2_62_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_62_pytest.log
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default norot default norot; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  echo "== $v"; timeout 600 python profiles/scripts/e2e_phases2.py 2>&1 | tail -11 | grep -E "total|T5|cross K|prefill" | tee -a gpurun_out/r2_62_phases_$v.txt
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
