#!/bin/bash
# round 2, call 78 (1 GPU): the allocation-free entry point's test ALONE in a fresh process, final library
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_78_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_78_pytest.log
