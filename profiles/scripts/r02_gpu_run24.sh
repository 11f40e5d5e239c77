#!/bin/bash
# round 2, call 24 (1 GPU): LATER dependent release, same box:
#   varAL fc1 (umma_act) releases fc2 only when its MMAs are done      varGL every lean split-K GEMM releases at MMA done
#   varT1 self-attention releases out_proj after its K pass            varT2 ... only at exit
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
for v in default varAL varGL varT1 varT2 default; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_24_bench_$v.json 2> gpurun_out/r2_24_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_24_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
