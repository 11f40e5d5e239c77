#!/bin/bash
# round 2, call 49 (1 GPU): full GPU suite, smoke(), default bench line (v18) and the B=128 line after the fc1 change
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_49_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_49_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_49_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_49_smoke.log
timeout 900 python bench.py > gpurun_out/r2_49_bench_bf16.json 2> gpurun_out/r2_49_bench_bf16.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/r2_49_bench_bf16.json
timeout 900 python bench.py --batch 128 --skip-cpu > gpurun_out/r2_49_bench_b128.json 2> gpurun_out/r2_49_bench_b128.err; echo "bench b128 rc=$?"; cut -c1-300 gpurun_out/r2_49_bench_b128.json
