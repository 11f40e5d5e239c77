#!/bin/bash
# round 2, call 63 (1 GPU): ragged description masks - mask words requested before the Wo prefetches and looked at after
# them (default) against looked at immediately (xbv1), same box; masked parity tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_generate.py tests/test_gpu_parity.py -m gpu -x -q -k "mask or ragged or cross or guidance" > gpurun_out/r2_63_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_63_pytest.log
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default xbv1 default xbv1; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  echo "== $v"; timeout 600 python profiles/scripts/loop_in_generate.py ragged 2>&1 | tail -3 | tee -a gpurun_out/r2_63_ragged_$v.txt
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
