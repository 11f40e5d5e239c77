#!/bin/bash
# round 2, call 45 (1 GPU): fc1 + activation with 32 batch rows per CTA and no X ring (default) against the 64-row / ring
# version (base), same box, B = 64 and B = 128
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "activation_gemm or heads_partials or several_decode" > gpurun_out/r2_45_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_45_pytest.log
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
for v in default base default base; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_45_bench_$v.json 2> gpurun_out/r2_45_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_45_bench_$v.json
done
for v in default base; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 --batch 128 > gpurun_out/r2_45_bench_b128_$v.json 2> gpurun_out/r2_45_bench_b128_$v.err; echo "b128 $v rc=$?"; line gpurun_out/r2_45_bench_b128_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
