#!/bin/bash
# round 2, call 47 (1 GPU): codec decode beside the token loop with a CTA limit on the persistent conv kernels and the
# step graph captured at high stream priority - whole generate() call against decode-after-the-loop, same box
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python profiles/scripts/overlap_time.py 128:0 128:32 128:48 128:64 256:48 256:64 256:96 > gpurun_out/r2_47_overlap.txt 2> gpurun_out/r2_47_overlap.err; echo "overlap rc=$?"; cat gpurun_out/r2_47_overlap.txt; tail -3 gpurun_out/r2_47_overlap.err
