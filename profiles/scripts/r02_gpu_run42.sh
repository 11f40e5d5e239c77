#!/bin/bash
# round 2, call 42 (1 GPU): per-phase SM-clock stamps of the cross-attention block inside the replayed graph (diagnostic
# library), after the single-lane TMA issue
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
PTTS_LIBRARY=$PWD/parler_tts_b200/libparler_b200_diag.so timeout 600 python profiles/scripts/xblock_stamps.py > gpurun_out/r2_42_xblock_phases.txt 2>&1; tail -9 gpurun_out/r2_42_xblock_phases.txt
