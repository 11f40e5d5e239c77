#!/bin/bash
# round 2, call 57 (1 GPU): the decode loop as 1 / 2 / 4 independent sub-batches on their own streams
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python profiles/scripts/two_streams.py > gpurun_out/r2_57_two_streams.txt 2>&1; echo "rc=$?"; tail -8 gpurun_out/r2_57_two_streams.txt
