#!/bin/bash
# round 2, call 13 (1 GPU): streaming generate() - GPU tests, then time-to-first-audio / total time at the benched size
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_streaming.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_13_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_13_pytest.log
tail -25 gpurun_out/r2_13_pytest.log
timeout 900 python profiles/scripts/stream_time.py 64 128 256 > gpurun_out/r2_13_stream_time.txt 2>&1; tail -12 gpurun_out/r2_13_stream_time.txt
