# pre-wait K staging in shared memory (fused self-attention): 0 / 64 / 128 rows; parity tests of the new path
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r6_pytest.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/r6_pytest.log
B="python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r6_$name.json 2> gpurun_out/r6_$name.err; echo "$name exit $?"; }
run stage0 PTTS_ATTN_STAGE=0
run stage64 PTTS_ATTN_STAGE=64
run stage128 PTTS_ATTN_STAGE=128
for f in stage0 stage64 stage128; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r6_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["value"]), "frames/s", round(d["ms_per_decode_step"], 4), "ms/step")
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/r6_$f.err").read()[-1500:])
PY
done
