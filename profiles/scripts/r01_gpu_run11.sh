# fused cross-attention block, second pass: out_proj releases it late, LayerNorm loads in one batch, phase stamps
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced or graph_replays or long_context" > gpurun_out/r11_pytest_new.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r11_pytest_new.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r11_$name.json 2> gpurun_out/r11_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r11_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r11_$name.err").read()[-1500:])
PY
}
run xb_late X=1
run xb_late_t16 PTTS_SPLIT_TARGET=16
run xb_early_t16 PTTS_XB_LATE=0 PTTS_SPLIT_TARGET=16
run xb_early PTTS_XB_LATE=0
run noxb PTTS_XBLOCK=0
run xb_late_nopf PTTS_L2PF=0
run xb_late_t16_nopf PTTS_SPLIT_TARGET=16 PTTS_L2PF=0
run xb_late_pct25 PTTS_L2PF_PCT=25
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r11_timeline.txt 2>&1; echo "timeline exit $?"
tail -25 gpurun_out/r11_timeline.txt
timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r11_xb_stamps.txt 2>&1; echo "stamps exit $?"
tail -12 gpurun_out/r11_xb_stamps.txt
PTTS_SPLIT_TARGET=16 timeout 300 python profiles/scripts/xblock_stamps.py > gpurun_out/r11_xb_stamps_t16.txt 2>&1; echo "stamps exit $?"
tail -12 gpurun_out/r11_xb_stamps_t16.txt
