#!/bin/bash
# round 2, call 33 (1 GPU): step epilogue launched with the programmatic-serialization attribute (resident when the
# lm_heads GEMM finishes; default) against the plain launch (varE0 = the library of the commit before), same box
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py tests/test_gpu_fast_kernels.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_33_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_33_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varE0 default varE0; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_33_bench_$v.json 2> gpurun_out/r2_33_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_33_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
