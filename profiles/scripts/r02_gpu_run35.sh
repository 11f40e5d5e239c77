#!/bin/bash
# round 2, call 35 (1 GPU): self-attention requests its first V rows under the LAST K batch (default) against the request
# after the K pass (varV0 = the library of the commit before), bf16 and fp8 KV cache, same box; attention tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_api.py tests/test_zz_no_alloc.py tests/test_gpu_baseline_configs.py -m gpu -x -q > gpurun_out/r2_35_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_35_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "attn us", round(d["roofline"]["dominant_kernel"]["ms_per_launch"]*1e3,2), "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varV0 default varV0; do for kv in bf16 fp8; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e --steps 2 --kv-dtype $kv > gpurun_out/r2_35_bench_${v}_$kv.json 2> gpurun_out/r2_35_bench_${v}_$kv.err; echo "$v $kv rc=$?"; line gpurun_out/r2_35_bench_${v}_$kv.json
done; done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
