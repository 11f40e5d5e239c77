# release-timing / cluster knobs once more under the lean kernels
set -x
mkdir -p gpurun_out
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 200 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r37_$name.json 2> gpurun_out/r37_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r37_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r37_$name.err").read()[-1500:])
PY
}
run default X=1
run act_late PTTS_ACT_EARLY=0
run xblate2 PTTS_XB_LATE=2
run cs2 PTTS_XB_CLUSTER=2
