#!/bin/bash
# round 2, call 43 (1 GPU): flakiness check - the full GPU suite three times in fresh processes, once in random-ish order
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for i in 1 2 3; do
  timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_43_pytest_$i.log 2>&1; echo "run $i rc=$?"; tail -2 gpurun_out/r2_43_pytest_$i.log
done
timeout 1800 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_streaming.py tests/test_gpu_generate.py -m gpu -x -q > gpurun_out/r2_43_pytest_alone.log 2>&1; echo "subset rc=$?"; tail -2 gpurun_out/r2_43_pytest_alone.log
