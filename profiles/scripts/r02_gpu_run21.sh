#!/bin/bash
# round 2, call 21 (1 GPU): plain 128 x 128 tile GEMM for M >= 256 (fast_gemm_big.cu) in the prefill / cross-KV / T5
# chains: GEMM and chain tests, kernel tables and the phase split of generate(), one bench line with its in-run check
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_baseline_configs.py tests/test_gpu_t5.py tests/test_gpu_parity.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_21_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_21_pytest.log
tail -12 gpurun_out/r2_21_pytest.log
timeout 600 python profiles/scripts/prefill_profile.py > gpurun_out/r2_21_prefill_profile.txt 2>&1; grep -v "^-" gpurun_out/r2_21_prefill_profile.txt | cut -c1-72,150-260 | grep "ptts\|====" | head -24
timeout 600 python profiles/scripts/e2e_phases2.py > gpurun_out/r2_21_e2e_phases2.txt 2>&1; tail -12 gpurun_out/r2_21_e2e_phases2.txt
timeout 900 python bench.py --skip-cpu > gpurun_out/r2_21_bench.json 2> gpurun_out/r2_21_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r2_21_bench.json").read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "e2e", d["e2e"] and round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],3), "check", d["check"])
except Exception as e:
    print("   no line:", e)
PY
