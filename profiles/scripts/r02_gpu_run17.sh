#!/bin/bash
# round 2, call 17 (1 GPU): persistent warp-specialised flash-decode with single-lane (elect.sync) issue and spinning
# test_wait in its producer / MMA-issuer lanes: unit tests, isolated timing, decode step with --self-attn-kernel 2
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention_ws.py -m gpu -x -q > gpurun_out/r2_17_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_17_pytest.log
tail -5 gpurun_out/r2_17_pytest.log
timeout 600 python profiles/scripts/attn_ws_time.py 130 466 898 > gpurun_out/r2_17_attn_ws_time.txt 2>&1; tail -14 gpurun_out/r2_17_attn_ws_time.txt
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],3),
          "check", d["check"].get("ok"), d["check"].get("codes_sha16"), d["check"].get("simple_kernels_logit_rel_err"))
except Exception as e:
    print("   no line:", e)
PY
}
for v in 0 2; do for kv in bf16 fp8; do
  timeout 900 python bench.py --skip-cpu --skip-e2e --self-attn-kernel $v --kv-dtype $kv > gpurun_out/r2_17_bench_sak${v}_$kv.json 2> gpurun_out/r2_17_bench_sak${v}_$kv.err; echo "sak $v $kv rc=$?"; line gpurun_out/r2_17_bench_sak${v}_$kv.json
done; done
