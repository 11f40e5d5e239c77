# out_proj releases the cross-attention block at MMA-done; per-k-block X barriers with the expectations issued first
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced or graph_replays or skinny" > gpurun_out/r18_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r18_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r18_$name.json 2> gpurun_out/r18_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r18_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r18_$name.err").read()[-1500:])
PY
}
run default X=1
run xblate2 PTTS_XB_LATE=2
run xbar0 PTTS_GEMM_XBAR=0
run xblate2_xbar0 PTTS_XB_LATE=2 PTTS_GEMM_XBAR=0
run default_again X=2
run xblate0 PTTS_XB_LATE=0
PTTS_XB_LATE=2 timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r18_timeline_late2.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r18_timeline_late2.txt
