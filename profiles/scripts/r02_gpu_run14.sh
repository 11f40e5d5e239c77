#!/bin/bash
# round 2, call 14 (1 GPU): streaming tests; same-box A/B of two variant libraries against the default build:
#   varA  -DPTTS_GEMM_XSPLIT=1      X slab as two box copies, MMAs of the first half under the second half's landing
#   varB  -DPTTS_ATTN_SPLIT_BIG=2   two KV splits per (utterance, head) at B=64 (2048 CTAs, two waves) + combine kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_streaming.py -m gpu -x -q > gpurun_out/r2_14_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_14_pytest.log
tail -5 gpurun_out/r2_14_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],3),
          "attn us", round(d["roofline"]["dominant_kernel"]["ms_per_launch"]*1e3,2), "check", d["check"].get("ok"), d["check"].get("codes_sha16"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varA varB default varA varB; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e > gpurun_out/r2_14_bench_$v.json 2> gpurun_out/r2_14_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_14_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
