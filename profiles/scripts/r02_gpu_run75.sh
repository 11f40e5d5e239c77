#!/bin/bash
# round 2, call 75 (1 GPU): cached decode steps (masked / unmasked) against the reference fixtures' own step logits, second decoder shape
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_api.py -m gpu -x -q -k "cached_steps or static_cache or second_decoder_shape" > gpurun_out/r2_75_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2_75_pytest.log
