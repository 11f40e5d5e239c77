#!/bin/bash
# round 2, call 27 (1 GPU): state of record v15 - full GPU suite, smoke(), default bench line (with the CPU baseline leg),
# reference arm, the other BASELINE configs, streaming timing
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_27_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_27_pytest.log
tail -4 gpurun_out/r2_27_pytest.log
timeout 900 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_27_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2_27_smoke.log
timeout 1200 python bench.py > gpurun_out/r2_27_bench.json 2> gpurun_out/r2_27_bench.err; echo "bench rc=$?"
timeout 1200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_27_bench_reference_arm.json 2> gpurun_out/r2_27_bench_reference_arm.err; echo "reference arm rc=$?"; tail -c 400 gpurun_out/r2_27_bench_reference_arm.json
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
line gpurun_out/r2_27_bench.json
run1() { name=$1; shift; timeout 900 python bench.py --skip-cpu "$@" > gpurun_out/r2_27_bench_$name.json 2> gpurun_out/r2_27_bench_$name.err; echo "$name rc=$?"; line gpurun_out/r2_27_bench_$name.json; }
run1 fp8 --kv-dtype fp8
run1 b256 --batch 256 --steps 2 --warmup 3 --e2e-steps 2
run1 b128 --batch 128 --steps 2 --warmup 3 --e2e-steps 2
run1 scaled_b32 --arch scaled --batch 32 --steps 2 --warmup 3 --e2e-steps 2
run1 long_b16 --batch 16 --frames 2600 --steps 2 --warmup 3 --e2e-steps 2
run1 sample --do-sample --steps 2 --warmup 3 --e2e-steps 2
timeout 900 python profiles/scripts/stream_time.py 64 128 > gpurun_out/r2_27_stream_time.txt 2>&1; tail -6 gpurun_out/r2_27_stream_time.txt
