#!/bin/bash
# round 2, call 53 (1 GPU): the loop inside generate() with and without the codec phase behind it, SM clocks sampled
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python profiles/scripts/loop_in_generate.py > gpurun_out/r2_53_loop_in_generate.txt 2>&1; echo "rc=$?"; tail -10 gpurun_out/r2_53_loop_in_generate.txt
