# GEMM: test_wait barriers (per k-block or single), split target 16; xblock: tensor-core scores, butterfly reduce;
# LayerNorm: 20 slice loads in flight
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r14_pytest.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/r14_pytest.log | cut -c1-300
PTTS_XB_CLUSTER=1 timeout 600 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini" > gpurun_out/r14_pytest_cs1.log 2>&1; echo "pytest cs1 exit $?"
tail -3 gpurun_out/r14_pytest_cs1.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r14_$name.json 2> gpurun_out/r14_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r14_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r14_$name.err").read()[-1500:])
PY
}
run cs2 X=1
run cs1 PTTS_XB_CLUSTER=1
run cs1_xbar PTTS_XB_CLUSTER=1 PTTS_GEMM_XBAR=1
run noxb PTTS_XBLOCK=0
run noxb_t30 PTTS_XBLOCK=0 PTTS_SPLIT_TARGET=30
run cs1_t30 PTTS_XB_CLUSTER=1 PTTS_SPLIT_TARGET=30
run cs1_early PTTS_XB_CLUSTER=1 PTTS_XB_LATE=0
run cs1_pdllate PTTS_XB_CLUSTER=1 PTTS_PDL_EARLY=0
PTTS_XB_CLUSTER=1 timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r14_timeline.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r14_timeline.txt
PTTS_XB_CLUSTER=1 timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r14_xb_stamps_cs1.txt 2>&1; echo "stamps exit $?"
tail -9 gpurun_out/r14_xb_stamps_cs1.txt
PTTS_XB_CLUSTER=1 timeout 300 python profiles/scripts/gemm_agg.py > gpurun_out/r14_gemm_agg.txt 2>&1; echo "agg exit $?"
tail -8 gpurun_out/r14_gemm_agg.txt
