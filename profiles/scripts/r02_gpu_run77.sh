#!/bin/bash
# round 2, call 77 (1 GPU): DAC rates trial incl. the tcgen05 conv path with permuted rates; full GPU suite on the final tree
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 60 python profiles/scripts/dac_rates_trial.py > gpurun_out/r2_77_dac_rates.txt 2>&1; echo "trial rc=$?"; tail -4 gpurun_out/r2_77_dac_rates.txt
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_77_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_77_pytest.log
tail -4 gpurun_out/r2_77_pytest.log
