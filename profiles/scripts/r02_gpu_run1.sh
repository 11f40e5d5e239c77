#!/bin/bash
# round 2, call 1: GPU test suite, smoke, bench lines (bf16 and fp8 KV cache)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r2_1_gpu.txt 2>&1
nproc >> gpurun_out/r2_1_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r2_1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_1_pytest.log
tail -5 gpurun_out/r2_1_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2_1_smoke.log
tail -3 gpurun_out/r2_1_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_1_bench.json 2> gpurun_out/r2_1_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --steps 3 --warmup 3 --kv-dtype fp8 --skip-cpu > gpurun_out/r2_1_bench_fp8.json 2> gpurun_out/r2_1_bench_fp8.err; echo "bench fp8 rc=$?"
tail -c 600 gpurun_out/r2_1_bench.err gpurun_out/r2_1_bench_fp8.err
