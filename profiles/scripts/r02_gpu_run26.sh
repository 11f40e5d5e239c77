#!/bin/bash
# round 2, call 26 (1 GPU): marginal in-graph cost of each kernel class (diagnostic library, PTTS_SKIP masks)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python profiles/scripts/ablate2.py > gpurun_out/r2_26_ablation.txt 2>&1; cat gpurun_out/r2_26_ablation.txt | tail -14
