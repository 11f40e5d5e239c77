#!/bin/bash
# round 2, call 25 (1 GPU): several decode steps per CUDA-graph replay (graph-to-graph hand-over removed between them)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "launches", d["gpu_launches"], "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
except Exception as e:
    print("   no line:", e)
PY
}
for u in 1 2 4 8 16 1; do
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 --steps-per-graph $u > gpurun_out/r2_25_bench_u$u.json 2> gpurun_out/r2_25_bench_u$u.err; echo "steps per graph $u rc=$?"; line gpurun_out/r2_25_bench_u$u.json
done
