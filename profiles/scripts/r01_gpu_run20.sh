# cross-attention block writes its per-head slices row-contiguously; the next LayerNorm pulls a row's 64 KB with one bulk copy
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r20_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/r20_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r20_$name.json 2> gpurun_out/r20_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r20_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r20_$name.err").read()[-1500:])
PY
}
run rowout X=1
run strided PTTS_XB_ROWOUT=0
run rowout_b X=2
run strided_b PTTS_XB_ROWOUT=0
run rowout_cs2 PTTS_XB_CLUSTER=2
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r20_timeline.txt 2>&1; echo "timeline exit $?"
tail -12 gpurun_out/r20_timeline.txt
