#!/bin/bash
# round 2, call 56 (2 GPUs): bench at N=2 the way the driver launches it, and the reference arm under torchrun
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2_56_bench_n2.json 2> gpurun_out/r2_56_bench_n2.err; echo "n2 rc=$?"; tail -1 gpurun_out/r2_56_bench_n2.json | cut -c1-500
timeout 600 python -m pytest tests -m gpu -x -q -k "nccl or sharded or two_rank or world" > gpurun_out/r2_56_pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_56_pytest_n2.log
