#!/bin/bash
# round 2, call 55 (1 GPU): ragged description masks - cross-attention block with the key-mask words fetched before the
# dependency wait (default) against the mask load inside the attention phase (xbprev), same box; masked parity tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_generate.py tests/test_gpu_parity.py -m gpu -x -q -k "mask or ragged or cross or guidance" > gpurun_out/r2_55_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_55_pytest.log
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default xbprev default xbprev; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  echo "== $v"; timeout 600 python profiles/scripts/loop_in_generate.py ragged 2>&1 | tail -5 | tee -a gpurun_out/r2_55_ragged_$v.txt
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
