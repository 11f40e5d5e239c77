# lean instantiations (optional paths compiled out) of the GEMM, the cross-attention block and the flash-decode kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r26_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/r26_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r26_$name.json 2> gpurun_out/r26_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r26_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r26_$name.err").read()[-1500:])
PY
}
run lean_all X=1
run full_all PTTS_GEMM_LEAN=0 PTTS_XB_LEAN=0 PTTS_ATTN_LEAN=0
run full_gemm PTTS_GEMM_LEAN=0
run full_xb PTTS_XB_LEAN=0
run full_attn PTTS_ATTN_LEAN=0
run lean_all_b X=2
