# fc1 pair finisher with the activation in a rolled loop over sums parked in shared memory
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q > gpurun_out/r23_pytest.log 2>&1; echo "pytest exit $?"
tail -6 gpurun_out/r23_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r23_$name.json 2> gpurun_out/r23_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r23_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r23_$name.err").read()[-1500:])
PY
}
run pair1 X=1
run pair0 PTTS_FC1_PAIR=0
run pair2 PTTS_FC1_PAIR=2
run pair1_b X=2
run pair0_b PTTS_FC1_PAIR=0

timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r23_timeline.txt 2>&1; echo "timeline exit $?"
tail -22 gpurun_out/r23_timeline.txt
