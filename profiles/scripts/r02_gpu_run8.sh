#!/bin/bash
# round 2, call 8 (kernel v2: independent MMA issuers, two softmax groups): warp-specialised persistent attention kernel - unit test, then the decode step with it (2) against the
# per-item kernels (1), both cache dtypes; every bench run carries its own token / logit check
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention_ws.py -m gpu -q -x > gpurun_out/r2_8_ws.log 2>&1
echo "ws unit rc=$?" | tee -a gpurun_out/r2_8_ws.log
tail -3 gpurun_out/r2_8_ws.log
for kv in bf16 fp8; do
  for k in 2; do
    timeout 600 python bench.py --steps 2 --warmup 2 --kv-dtype $kv --self-attn-kernel $k --skip-cpu --skip-e2e \
      > gpurun_out/r2_8_bench_${kv}_k${k}.json 2> gpurun_out/r2_8_bench_${kv}_k${k}.err
    echo "bench $kv kernel=$k rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_8_bench_${kv}_k${k}.json").read().strip().splitlines()[-1])
    dk=d["roofline"]["dominant_kernel"]
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "step frac", round(d["roofline"]["frac"],3),
          "check", d["check"].get("ok"), d["check"].get("simple_kernels_logit_rel_err"))
except Exception as e:
    print("  no line:", e)
PY
    tail -c 300 gpurun_out/r2_8_bench_${kv}_k${k}.err
  done
done
