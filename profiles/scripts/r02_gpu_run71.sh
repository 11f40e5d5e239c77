#!/bin/bash
# round 2, call 71 (1 GPU): test_gpu_parity.py after the per-decision (unconditional) token assertion of the 24-layer fp32 test
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_71_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_71_pytest.log
