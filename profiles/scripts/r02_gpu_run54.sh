#!/bin/bash
# round 2, call 54 (1 GPU): generate() drops all-ones attention masks before the decode loop - tests, the loop inside
# generate(), default bench line
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_54_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_54_pytest.log
timeout 600 python profiles/scripts/loop_in_generate.py > gpurun_out/r2_54_loop_in_generate.txt 2>&1; echo "rc=$?"; tail -8 gpurun_out/r2_54_loop_in_generate.txt
timeout 600 python profiles/scripts/e2e_phases2.py > gpurun_out/r2_54_e2e_phases.txt 2>&1; echo "phases rc=$?"; tail -11 gpurun_out/r2_54_e2e_phases.txt
timeout 900 python bench.py > gpurun_out/r2_54_bench_bf16.json 2> gpurun_out/r2_54_bench_bf16.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2_54_bench_bf16.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms/step", round(d["ms_per_decode_step"],4), "e2e", round(d["e2e"]["value"]), "check", d["check"]["ok"], d["check"]["codes_sha16"], "frac", round(d["roofline"]["frac"],4))
PY
