#!/bin/bash
# round 2, call 67 (1 GPU): state of record v21 (DAC rates read from the checkpoint; no kernel change since v20) - full
# GPU suite, smoke(), default bench line with the CPU baseline leg, launch list of one eager step of the same library
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_67_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_67_pytest.log
tail -4 gpurun_out/r2_67_pytest.log
timeout 900 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_67_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_67_smoke.log
timeout 1200 python bench.py > gpurun_out/r2_67_bench.json 2> gpurun_out/r2_67_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2_67_bench.json",):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "e2e", d["e2e"] and round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],3), "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
    except Exception as e:
        print("   no line:", e)
PY
timeout 600 python profiles/scripts/ncu_step.py bf16 > gpurun_out/r2_67_plain.log 2>&1 &&
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/r2_67_ncu_launches.csv \
  --metrics gpu__time_duration.sum python profiles/scripts/ncu_step.py bf16 > gpurun_out/r2_67_ncu.log 2>&1; echo "ncu launch list rc=$?"
