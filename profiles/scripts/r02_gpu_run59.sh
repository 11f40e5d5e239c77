#!/bin/bash
# round 2, call 59 (1 GPU): sampling epilogue with the ordered keys in registers and the exponentials taken by all lanes
# (default) against the previous epilogue (epiprev): sampling parity tests, top-k 50 bench lines on one box
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_generate.py tests/test_gpu_fast_kernels.py -m gpu -x -q > gpurun_out/r2_59_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_59_pytest.log
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
for v in default epiprev default epiprev; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 --do-sample > gpurun_out/r2_59_bench_sample_$v.json 2> gpurun_out/r2_59_bench_sample_$v.err; echo "$v rc=$?"; line gpurun_out/r2_59_bench_sample_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 > gpurun_out/r2_59_bench_greedy.json 2> gpurun_out/r2_59_bench_greedy.err; echo "greedy rc=$?"; line gpurun_out/r2_59_bench_greedy.json
