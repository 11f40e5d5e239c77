#!/bin/bash
# round 2, call 11 (1 GPU): full GPU suite on the rebuilt library, default bench line, single-GPU lines of the other
# BASELINE configs (C3 per-GPU batches 256 / 128, C4 scaled decoder B=32, C5 long-form B=16 x 2600 frames, top-k
# sampling wording of configs[1]), end-to-end phase split
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_11_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_11_pytest.log
tail -3 gpurun_out/r2_11_pytest.log
timeout 900 python bench.py > gpurun_out/r2_11_bench.json 2> gpurun_out/r2_11_bench.err; echo "bench rc=$?"
run1() { # name, args...
  name=$1; shift
  timeout 900 python bench.py --skip-cpu "$@" > gpurun_out/r2_11_bench_$name.json 2> gpurun_out/r2_11_bench_$name.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_11_bench_$name.json").read().strip().splitlines()[-1])
    print("  ", d["config"]["workload"][:70], "| ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]),
          "e2e", d["e2e"] and round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],3), "check", d["check"].get("ok"))
except Exception as e:
    print("   no line:", e)
PY
}
tail -c 600 gpurun_out/r2_11_bench.json
run1 b256 --batch 256 --steps 2 --warmup 3 --e2e-steps 2
run1 b128 --batch 128 --steps 2 --warmup 3 --e2e-steps 2
run1 scaled_b32 --arch scaled --batch 32 --steps 2 --warmup 3 --e2e-steps 2
run1 long_b16 --batch 16 --frames 2600 --steps 2 --warmup 3 --e2e-steps 2
run1 sample --do-sample --steps 2 --warmup 3 --e2e-steps 2
run1 fp8 --kv-dtype fp8 --steps 3 --warmup 3
timeout 600 python profiles/scripts/e2e_phases.py > gpurun_out/r2_11_e2e_phases.txt 2>&1; tail -15 gpurun_out/r2_11_e2e_phases.txt
