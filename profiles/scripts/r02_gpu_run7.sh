#!/bin/bash
# round 2, call 7: the no-alloc entry point's test alone (fresh process), then the whole GPU suite with -x (it runs last)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2_7_noalloc_alone.log 2>&1
echo "alone rc=$?" | tee -a gpurun_out/r2_7_noalloc_alone.log
tail -15 gpurun_out/r2_7_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r2_7_pytest_all.log 2>&1
echo "all rc=$?" | tee -a gpurun_out/r2_7_pytest_all.log
tail -8 gpurun_out/r2_7_pytest_all.log
