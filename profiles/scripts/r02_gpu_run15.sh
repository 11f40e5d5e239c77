#!/bin/bash
# round 2, call 15 (1 GPU): default build (X slab in 2 parts) against varX4 (4 parts) and varB (attention: two KV splits
# per (utterance, head) at B=64 + combine kernel; --skip-check because the eager check pass plans one split on its short
# cache); finer end-to-end phase split
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],3),
          "attn us", round(d["roofline"]["dominant_kernel"]["ms_per_launch"]*1e3,2), "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varX4 varB default varX4 varB; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  extra=""; [ $v = varB ] && extra="--skip-check"
  timeout 900 python bench.py --skip-cpu --skip-e2e $extra > gpurun_out/r2_15_bench_$v.json 2> gpurun_out/r2_15_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_15_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
timeout 600 python profiles/scripts/e2e_phases2.py > gpurun_out/r2_15_e2e_phases2.txt 2>&1; tail -14 gpurun_out/r2_15_e2e_phases2.txt
