#!/bin/bash
# round 2, call 37 (1 GPU): DAC conv epilogue pulls the residual pieces of all its column groups into L2 while the tile's
# MMAs still run (default) against one group of residual loads in flight (varR0 = the library before), same box
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_generate.py tests/test_gpu_fast_kernels.py -m gpu -x -q -k "dac or DAC or decode or generate" > gpurun_out/r2_37_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_37_pytest.log
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varR0 default varR0; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  echo "== $v"; timeout 600 python profiles/scripts/dac_time.py 32 860 2>&1 | grep "dac fast"; timeout 600 python profiles/scripts/dac_time.py 32 860 2>&1 | grep "dac fast"
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
timeout 600 python profiles/scripts/dac_layers.py > gpurun_out/r2_37_dac_layers.txt 2>&1; grep "k/s= 1\|total" gpurun_out/r2_37_dac_layers.txt
