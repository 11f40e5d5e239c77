# xblock v3 (LayerNorm shared by clusters of 1 / 2 / 4 heads, K/V in registers) + per-k-block X barriers in the TMA GEMM
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced or graph_replays or long_context or skinny" > gpurun_out/r13_pytest.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r13_pytest.log | cut -c1-300
for cs in 1 4; do
PTTS_XB_CLUSTER=$cs timeout 600 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini" > gpurun_out/r13_pytest_cs$cs.log 2>&1; echo "pytest cs$cs exit $?"
tail -3 gpurun_out/r13_pytest_cs$cs.log | cut -c1-300
done
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r13_$name.json 2> gpurun_out/r13_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r13_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r13_$name.err").read()[-1500:])
PY
}
run cs2 X=1
run cs1 PTTS_XB_CLUSTER=1
run cs4 PTTS_XB_CLUSTER=4
run noxb PTTS_XBLOCK=0
run cs2_kb4 PTTS_KBMAX=4
run noxb_kb4 PTTS_XBLOCK=0 PTTS_KBMAX=4
run cs2_early PTTS_XB_LATE=0
run cs1_t16 PTTS_XB_CLUSTER=1 PTTS_SPLIT_TARGET=16
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r13_timeline.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r13_timeline.txt
timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r13_xb_stamps.txt 2>&1; echo "stamps exit $?"
tail -9 gpurun_out/r13_xb_stamps.txt
PTTS_XB_CLUSTER=1 timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r13_xb_stamps_cs1.txt 2>&1; echo "stamps exit $?"
tail -9 gpurun_out/r13_xb_stamps_cs1.txt
timeout 300 python profiles/scripts/gemm_agg.py > gpurun_out/r13_gemm_agg.txt 2>&1; echo "agg exit $?"
tail -8 gpurun_out/r13_gemm_agg.txt
