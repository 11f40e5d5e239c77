#!/bin/bash
# round 2, call 38 (1 GPU): LayerNorm folds its row-contiguous slices (the 16 per-head slices of the cross-attention block) with
# eight shared-memory loads in flight (default) against the rolled loop (varL0), same box; also the sanity run after the
# removal of the variant macros
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_38_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_38_pytest.log
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
for v in default varL0 default varL0; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_38_bench_$v.json 2> gpurun_out/r2_38_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_38_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
