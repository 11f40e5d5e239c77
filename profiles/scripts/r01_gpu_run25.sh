# does the (unused) pair-finisher code in the GEMM kernel cost anything?  same box, two builds of the library
set -x
mkdir -p gpurun_out
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r25_$name.json 2> gpurun_out/r25_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r25_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r25_$name.err").read()[-1500:])
PY
}
run with_pair_code X=1
run no_pair_code PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_nopair.so
run with_pair_code_b X=2
run no_pair_code_b PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_nopair.so
