# same-box A/B of three builds: previous commit (lean GEMM / block / attention), + rolled TMA loops and one store loop in
# the GEMM, + lean LayerNorm / reduce+activation
set -x
mkdir -p gpurun_out
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r28_$name.json 2> gpurun_out/r28_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r28_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r28_$name.err").read()[-1500:])
PY
}
run prev PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_prev.so
run gemmonly PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_gemmonly.so
run cur X=1
run prev_b PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_prev.so
run gemmonly_b PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_gemmonly.so
run cur_b X=2
