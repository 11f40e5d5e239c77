#!/bin/bash
# round 2, call 73 (1 GPU): full GPU suite + smoke() on the final tree of the round (tests and host code changed since call 67,
# kernels did not)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r2_73_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_73_pytest.log
tail -4 gpurun_out/r2_73_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_73_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_73_smoke.log
