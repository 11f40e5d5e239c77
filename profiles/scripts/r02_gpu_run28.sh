#!/bin/bash
# round 2, call 28 (1 GPU): 32-byte global accesses (LDG.256 / STG.256), same box:
#   default: DAC conv epilogue moves its fp32 rows as 2 x 32 bytes per lane (residual loads and stores)
#   varD0:   the same with 4 x 16 bytes (-DPTTS_DAC_WIDE=0)
#   varW:    self-attention with 32-byte K / V loads (4 threads per key row, -DPTTS_ATTN_WIDE=1)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_generate.py tests/test_gpu_fast_kernels.py -m gpu -x -q > gpurun_out/r2_28_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_28_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "check", d["check"].get("ok"), d["check"].get("codes_sha16"), d["check"].get("simple_kernels_logit_rel_err"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varD0 default varD0; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  echo "== $v"; timeout 600 python profiles/scripts/dac_time.py 32 860 2>&1 | grep "dac fast"
done
for v in default varW default varW; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_28_bench_$v.json 2> gpurun_out/r2_28_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_28_bench_$v.json
done
cp parler_tts_b200/libparler_b200_varW.so parler_tts_b200/libparler_b200.so
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_baseline_configs.py -m gpu -x -q > gpurun_out/r2_28_pytest_wide.log 2>&1; echo "pytest (wide attention) rc=$?"; tail -3 gpurun_out/r2_28_pytest_wide.log
cp /tmp/default.so parler_tts_b200/libparler_b200.so
