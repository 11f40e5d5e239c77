#!/bin/bash
# round 2, call 18 (1 GPU): lm_heads as split-K partial sums folded by the epilogue kernel (default) against the STORE GEMM
# + logits matrix (varH0 = -DPTTS_HEADS_PARTIAL=0), same box; tests of the fused step against the step-at-a-time loop
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py tests/test_gpu_api.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_18_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_18_pytest.log
tail -12 gpurun_out/r2_18_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],3),
          "check", d["check"].get("ok"), d["check"].get("codes_sha16"), d["check"].get("eager_replay_token_match"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varH0 default varH0; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e > gpurun_out/r2_18_bench_$v.json 2> gpurun_out/r2_18_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_18_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
timeout 900 python bench.py --skip-cpu --skip-e2e --do-sample > gpurun_out/r2_18_bench_sample.json 2> gpurun_out/r2_18_bench_sample.err; echo "sample rc=$?"; line gpurun_out/r2_18_bench_sample.json
