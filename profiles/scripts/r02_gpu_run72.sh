#!/bin/bash
# round 2, call 72 (1 GPU): is the flash-decode kernel limited by HBM or by the path from L2 into the SMs?  ptts_attention over
# 24 distinct caches (HBM) against one cache launched back to back (L2 when K+V fit), current library
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 120 python profiles/scripts/attn_l2.py > gpurun_out/r2_72_attn_l2.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2_72_attn_l2.txt
