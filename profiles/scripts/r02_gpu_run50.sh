#!/bin/bash
# round 2, call 50 (1 GPU): 128 x 128 tile GEMM with the row-wise (coalesced) epilogue - tests, kernel table of T5 / prefill,
# phases of generate()
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_t5.py tests/test_gpu_parity.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_50_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_50_pytest.log
timeout 600 python profiles/scripts/prefill_profile.py > gpurun_out/r2_50_prefill_kernels.txt 2>&1; echo "prefill_profile rc=$?"; grep -E "gemm_big|attn_prefill|t5_attention|Self CUDA time" gpurun_out/r2_50_prefill_kernels.txt | cut -c1-60,130-220
timeout 600 python profiles/scripts/e2e_phases2.py > gpurun_out/r2_50_e2e_phases.txt 2>&1; echo "phases rc=$?"; tail -12 gpurun_out/r2_50_e2e_phases.txt
