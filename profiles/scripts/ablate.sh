#!/bin/bash
# Marginal cost of each kernel class inside the replayed decode-step graph (timing only; outputs are garbage when a
# class is skipped).  Usage: ablate.sh [extra bench args]
for gemm in cluster ring; do
for skip in 0 1 2 4 8 16 32 64 255; do
  PTTS_GEMM=$gemm PTTS_SKIP=$skip python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu "$@" 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('gemm=$gemm skip=$skip ms_per_decode_step=%.4f launches/step=%d' % (d['ms_per_decode_step'], d['gpu_launches'] / 2 / 868))"
done; done
