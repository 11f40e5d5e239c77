#!/bin/bash
# round 2, call 68 (1 GPU): the other BASELINE configs and the opt-in / sampling lines on the v21 library
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ", d["config"]["workload"][:64], "| ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]),
          "e2e", d["e2e"] and round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],3), "check", d["check"].get("ok"))
except Exception as e:
    print("   no line:", e)
PY
}
run1() { name=$1; shift; timeout 600 python bench.py --skip-cpu "$@" > gpurun_out/r2_68_bench_$name.json 2> gpurun_out/r2_68_bench_$name.err; echo "$name rc=$?"; line gpurun_out/r2_68_bench_$name.json; }
run1 fp8kv --kv-dtype fp8
run1 sample --do-sample --steps 2 --warmup 3 --e2e-steps 3
run1 b128 --batch 128 --steps 2 --warmup 3 --e2e-steps 2
run1 b256 --batch 256 --steps 2 --warmup 3 --e2e-steps 2
run1 scaled_b32 --arch scaled --batch 32 --steps 2 --warmup 3 --e2e-steps 2
run1 long_b16 --batch 16 --frames 2600 --steps 2 --warmup 3 --e2e-steps 2
