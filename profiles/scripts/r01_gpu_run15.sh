# xblock: TMA weight slabs, gamma/beta staged; GEMM: single X barrier default; PDL early/late re-measured
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r15_pytest.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/r15_pytest.log | cut -c1-300
PTTS_XB_CLUSTER=2 timeout 600 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini" > gpurun_out/r15_pytest_cs2.log 2>&1; echo "pytest cs2 exit $?"
tail -3 gpurun_out/r15_pytest_cs2.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r15_$name.json 2> gpurun_out/r15_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r15_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r15_$name.err").read()[-1500:])
PY
}
run default X=1
run pdllate PTTS_PDL_EARLY=0
run cs2 PTTS_XB_CLUSTER=2
run xb_early PTTS_XB_LATE=0
run noxb PTTS_XBLOCK=0
run noxb_pdllate PTTS_XBLOCK=0 PTTS_PDL_EARLY=0
run kb4 PTTS_KBMAX=4
run pf25 PTTS_XB_L2PF_PCT=25
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r15_timeline.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r15_timeline.txt
timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r15_xb_stamps.txt 2>&1; echo "stamps exit $?"
tail -9 gpurun_out/r15_xb_stamps.txt
timeout 300 python profiles/scripts/gemm_agg.py > gpurun_out/r15_gemm_agg.txt 2>&1; echo "agg exit $?"
tail -8 gpurun_out/r15_gemm_agg.txt
