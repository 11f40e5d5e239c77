#!/bin/bash
# round 2, call 30 (2 GPUs): multi-step graph test, the 2-rank NCCL sharding test and the N=2 bench line on the v15 library
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -q -x -k "several_decode_steps or graph_replays" > gpurun_out/r2_30_graph_test.log 2>&1; echo "graph tests rc=$?"; tail -3 gpurun_out/r2_30_graph_test.log
timeout 600 python -m pytest tests/test_gpu_sharded_nccl.py -m gpu -q -x > gpurun_out/r2_30_nccl_test.log 2>&1; echo "nccl test rc=$?"; tail -3 gpurun_out/r2_30_nccl_test.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2_30_bench_n2.json 2> gpurun_out/r2_30_bench_n2.err; echo "n2 rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r2_30_bench_n2.json").read().strip().splitlines()[-1])
    print("  n_gpus", d["n_gpus"], "ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "per gpu", round(d["value_per_gpu"]), "e2e", d["e2e"] and round(d["e2e"]["value"]), "check", d["check"].get("ok"))
except Exception as e:
    print("   no line:", e)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 \
  bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2_30_bench_n2_reference_arm.json 2> gpurun_out/r2_30_bench_n2_reference_arm.err; echo "n2 reference arm rc=$?"; tail -c 300 gpurun_out/r2_30_bench_n2_reference_arm.json
