#!/bin/bash
# round 2, call 48 (1 GPU): timeline of the chunk decodes beside the loop (when each one ran, when the loop ended)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
OVL_TIMELINE=1 OVL_PRIO=-1 timeout 900 python profiles/scripts/overlap_time.py 256:0 256:64 128:64 > gpurun_out/r2_48_overlap.txt 2> gpurun_out/r2_48_overlap.err; echo "overlap rc=$?"; cat gpurun_out/r2_48_overlap.txt; tail -3 gpurun_out/r2_48_overlap.err
