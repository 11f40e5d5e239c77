# new defaults (late releases, cluster 2): whole GPU suite, A/B around the defaults, timeline / stamps / GEMM phases
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r17_pytest_all.log 2>&1; echo "pytest-all exit $?"
tail -8 gpurun_out/r17_pytest_all.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r17_$name.json 2> gpurun_out/r17_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r17_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r17_$name.err").read()[-1500:])
PY
}
run default X=1
run cs1 PTTS_XB_CLUSTER=1
run gemm_early PTTS_PDL_EARLY=1
run xbout_early PTTS_XB_LATE=0
run t30 PTTS_SPLIT_TARGET=30
run noxb PTTS_XBLOCK=0
run l2pf0 PTTS_L2PF=0
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r17_timeline.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r17_timeline.txt
timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r17_xb_stamps.txt 2>&1; echo "stamps exit $?"
tail -9 gpurun_out/r17_xb_stamps.txt
timeout 300 python profiles/scripts/gemm_agg.py > gpurun_out/r17_gemm_agg.txt 2>&1; echo "agg exit $?"
tail -8 gpurun_out/r17_gemm_agg.txt
