# fused cross-attention block (fast_xblock.cu) + weight prefetch hints (PTTS_WPF): parity first, then A/B bench lines
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced or graph_replays or long_context" > gpurun_out/r10_pytest_new.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r10_pytest_new.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r10_$name.json 2> gpurun_out/r10_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r10_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r10_$name.err").read()[-1500:])
PY
}
run xb_wpf15 PTTS_WPF=15
run xb_wpf0 PTTS_WPF=0
run noxb_wpf0 PTTS_XBLOCK=0 PTTS_WPF=0
run noxb_wpf15 PTTS_XBLOCK=0 PTTS_WPF=15
run xb_wpf1 PTTS_WPF=1
run xb_wpf2 PTTS_WPF=2
run xb_wpf4 PTTS_WPF=4
run xb_wpf8 PTTS_WPF=8
run xb_wpf15_nopf PTTS_WPF=15 PTTS_L2PF=0
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r10_timeline.txt 2>&1; echo "timeline exit $?"
cat gpurun_out/r10_timeline.txt | tail -25
