#!/bin/bash
# round 2, call 76 (1 GPU): DAC decoder with other (even) up-sampling rates read from the checkpoint, against the oracle
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 100 python profiles/scripts/dac_rates_trial.py > gpurun_out/r2_76_dac_rates.txt 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2_76_dac_rates.txt
