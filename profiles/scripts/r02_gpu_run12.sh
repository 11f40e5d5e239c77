#!/bin/bash
# round 2, call 12 (1 GPU): single-lane (elect.sync) issue paths + second-generation lean split-K GEMM (3-D box copies,
# unrolled MMA issue).  Full GPU suite, then same-box A/B of the decode step and of the DAC decode against the
# previous commit's library (parler_tts_b200/libparler_b200_prev.so, built here from HEAD before the change).
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_12_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_12_pytest.log
tail -3 gpurun_out/r2_12_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "e2e", d["e2e"] and round(d["e2e"]["value"]),
          "frac", round(d["roofline"]["frac"],3), "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
except Exception as e:
    print("   no line:", e)
PY
}
for rep in 1 2; do
  timeout 900 python bench.py --skip-cpu > gpurun_out/r2_12_bench_new_$rep.json 2> gpurun_out/r2_12_bench_new_$rep.err; echo "new $rep rc=$?"; line gpurun_out/r2_12_bench_new_$rep.json
done
timeout 900 python bench.py --skip-cpu --kv-dtype fp8 > gpurun_out/r2_12_bench_new_fp8.json 2> gpurun_out/r2_12_bench_new_fp8.err; echo "new fp8 rc=$?"; line gpurun_out/r2_12_bench_new_fp8.json
timeout 600 python profiles/scripts/e2e_phases.py > gpurun_out/r2_12_e2e_phases_new.txt 2>&1; tail -2 gpurun_out/r2_12_e2e_phases_new.txt
timeout 900 python bench.py --skip-cpu --batch 16 --frames 2600 --steps 2 --warmup 3 --e2e-steps 2 > gpurun_out/r2_12_bench_long_b16.json 2> gpurun_out/r2_12_bench_long_b16.err; echo "long rc=$?"; line gpurun_out/r2_12_bench_long_b16.json
# previous library on the same box
cp parler_tts_b200/libparler_b200.so /tmp/new.so
cp parler_tts_b200/libparler_b200_prev.so parler_tts_b200/libparler_b200.so
timeout 900 python bench.py --skip-cpu > gpurun_out/r2_12_bench_prev.json 2> gpurun_out/r2_12_bench_prev.err; echo "prev rc=$?"; line gpurun_out/r2_12_bench_prev.json
timeout 600 python profiles/scripts/e2e_phases.py > gpurun_out/r2_12_e2e_phases_prev.txt 2>&1; tail -2 gpurun_out/r2_12_e2e_phases_prev.txt
cp /tmp/new.so parler_tts_b200/libparler_b200.so
