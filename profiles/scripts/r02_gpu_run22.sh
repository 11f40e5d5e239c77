#!/bin/bash
# round 2, call 22 (1 GPU): DAC decode batch chunking at B=64 x 860 frames
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python profiles/scripts/dac_chunking.py > gpurun_out/r2_22_dac_chunking.txt 2>&1; tail -8 gpurun_out/r2_22_dac_chunking.txt
