#!/bin/bash
# round 2, call 74 (1 GPU): delay-pattern kernels against 48 randomised reference cases (tests/golden/delay_patterns_random.npz)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "delay" > gpurun_out/r2_74_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_74_pytest.log
