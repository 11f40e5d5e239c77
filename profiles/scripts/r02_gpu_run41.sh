#!/bin/bash
# round 2, call 41 (1 GPU): unit test of the staged prefill attention against torch and the simple family
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "prefill_attention" > gpurun_out/r2_41_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2_41_pytest.log
