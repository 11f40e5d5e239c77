#!/bin/bash
# round 2, call 29 (1 GPU): varQE = only the QKV GEMM releases its dependent (the self-attention kernel) BEFORE its own
# wait: attention CTAs resident and their first K batch in flight while LayerNorm / QKV still run
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
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
for v in default varQE default varQE; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_29_bench_$v.json 2> gpurun_out/r2_29_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_29_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
