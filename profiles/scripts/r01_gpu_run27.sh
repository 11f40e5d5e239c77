# lean LayerNorm / reduce+activation instantiations, single predicated store loop in the GEMM epilogue
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r27_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/r27_pytest.log | cut -c1-300
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r27_$name.json 2> gpurun_out/r27_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r27_$name.json").read().strip().splitlines()[-1])
    print("$name", round(d["ms_per_decode_step"], 4), "ms/decode step", round(d["value"]), "frames/s")
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/r27_$name.err").read()[-1500:])
PY
}
run lean_all X=1
run lean_all_b X=2
run cs2 PTTS_XB_CLUSTER=2
run gemm_early PTTS_PDL_EARLY=1
run ln_early PTTS_LN_EARLY=1
run xb_early PTTS_XB_EARLY=1
run t30 PTTS_SPLIT_TARGET=30
run kb4 PTTS_KBMAX=4
