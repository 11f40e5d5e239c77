#!/bin/bash
# round 2, call 60 (1 GPU): state of record v20 - full GPU suite, smoke(), default line, top-k sampling line (with e2e)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_60_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_60_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_60_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_60_smoke.log
line() { python - "$1" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("  value", round(d["value"]), "ms/step", round(d.get("ms_per_decode_step",0),4), "e2e", round(d["e2e"]["value"]) if d.get("e2e") else None, "check", (d.get("check") or {}).get("ok"), "frac", round((d.get("roofline") or {}).get("frac",0),4))
PY
}
timeout 900 python bench.py > gpurun_out/r2_60_bench_bf16.json 2> gpurun_out/r2_60_bench_bf16.err; echo "bench rc=$?"; line gpurun_out/r2_60_bench_bf16.json
timeout 900 python bench.py --do-sample --skip-cpu > gpurun_out/r2_60_bench_sample.json 2> gpurun_out/r2_60_bench_sample.err; echo "bench sample rc=$?"; line gpurun_out/r2_60_bench_sample.json
