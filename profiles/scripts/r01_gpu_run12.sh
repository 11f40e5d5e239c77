# xblock v2 (cluster-shared LayerNorm, K/V in registers before the wait, K-split q GEMM) + TMA operand loads in the GEMM
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r12_pytest.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r12_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r12_$name.json 2> gpurun_out/r12_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r12_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r12_$name.err").read()[-1500:])
PY
}
run default X=1
run notma PTTS_GEMM_TMA=0
run noxb PTTS_XBLOCK=0
run noxb_notma PTTS_XBLOCK=0 PTTS_GEMM_TMA=0
run xb_early PTTS_XB_LATE=0
run xb_pf25 PTTS_XB_L2PF_PCT=25
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r12_timeline.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r12_timeline.txt
timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r12_xb_stamps.txt 2>&1; echo "stamps exit $?"
tail -12 gpurun_out/r12_xb_stamps.txt
timeout 300 python profiles/scripts/gemm_agg.py > gpurun_out/r12_gemm_agg.txt 2>&1; echo "agg exit $?"
tail -12 gpurun_out/r12_gemm_agg.txt
