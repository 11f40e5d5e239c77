#!/bin/bash
# round 2, call 52 (1 GPU): tile GEMM epilogue with 16 residual rows in flight - tests, launch list of T5 / prefill, phases
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_t5.py tests/test_gpu_parity.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_52_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_52_pytest.log
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2_52_prefill_launches.csv python profiles/scripts/prefill_ncu.py > gpurun_out/r2_52_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r2_52_ncu.log
python profiles/scripts/launch_table.py gpurun_out/r2_52_prefill_launches.csv > gpurun_out/r2_52_prefill_table.txt; cat gpurun_out/r2_52_prefill_table.txt
timeout 600 python profiles/scripts/e2e_phases2.py > gpurun_out/r2_52_e2e_phases.txt 2>&1; echo "phases rc=$?"; tail -11 gpurun_out/r2_52_e2e_phases.txt
