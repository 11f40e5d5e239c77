# who releases its dependent early? (with TMA slab loads the successors' bursts collide with the critical-path kernel)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced or graph_replays or layernorm" > gpurun_out/r16_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r16_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r16_$name.json 2> gpurun_out/r16_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r16_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r16_$name.err").read()[-1500:])
PY
}
run default X=1
run ln0 PTTS_LN_EARLY=0
run act0 PTTS_ACT_EARLY=0
run xb0 PTTS_XB_EARLY=0
run gemm0 PTTS_PDL_EARLY=0
run ln0_gemm0 PTTS_LN_EARLY=0 PTTS_PDL_EARLY=0
run ln0_act0_gemm0 PTTS_LN_EARLY=0 PTTS_ACT_EARLY=0 PTTS_PDL_EARLY=0
run all0 PTTS_LN_EARLY=0 PTTS_ACT_EARLY=0 PTTS_PDL_EARLY=0 PTTS_XB_EARLY=0
run cs2_ln0_gemm0 PTTS_XB_CLUSTER=2 PTTS_LN_EARLY=0 PTTS_PDL_EARLY=0
run cs2 PTTS_XB_CLUSTER=2
run ln0_xb0 PTTS_LN_EARLY=0 PTTS_XB_EARLY=0
