#!/bin/bash
# round 2, call 70 (2 GPUs): bench at N=2 on the v21 library, launched the way the driver launches it
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 140 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 --e2e-steps 2 --skip-cpu > gpurun_out/r2_70_bench_n2.json 2> gpurun_out/r2_70_bench_n2.err; echo "n2 rc=$?"; tail -1 gpurun_out/r2_70_bench_n2.json | cut -c1-400
