# A/B: dependent-kernel release before the GEMM's own wait (PTTS_PDL_EARLY), 64 KB W slabs per CTA (PTTS_KBMAX=4)
set -x
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r4_$name.json 2> gpurun_out/r4_$name.err; echo "$name exit $?"; }
run base PTTS_X=0
run early PTTS_PDL_EARLY=1
run kb4 PTTS_KBMAX=4
run early_kb4 PTTS_PDL_EARLY=1 PTTS_KBMAX=4
run pf2 PTTS_L2PF=2
PTTS_PDL_EARLY=1 PTTS_KBMAX=4 timeout 600 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r4_pytest_early.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/r4_pytest_early.log
for f in base early kb4 early_kb4 pf2; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r4_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["value"]), "frames/s", round(d["ms_per_decode_step"], 4), "ms/step")
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/r4_$f.err").read()[-1500:])
PY
done
