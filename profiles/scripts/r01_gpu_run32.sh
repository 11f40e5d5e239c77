# cross-attention block without mask tests and ragged-group predicates when neither is needed: same-box A/B + tests
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced" > gpurun_out/r32_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r32_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r32_$name.json 2> gpurun_out/r32_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r32_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r32_$name.err").read()[-1500:])
PY
}
run prev PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_prev.so
run plain X=1
run prev_b PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_prev.so
run plain_b X=2
