# sanity of the last commit (block falls back when tensor maps are unavailable): fast-kernel tests + one bench line
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fast_kernels.py -m gpu -x -q -k "cross_attention_block or scaled_and_mini or teacher_forced or graph_replays" > gpurun_out/r35_pytest.log 2>&1; echo "pytest exit $?"
tail -3 gpurun_out/r35_pytest.log | cut -c1-300
timeout 300 python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu > gpurun_out/r35_bench.json 2> gpurun_out/r35_bench.err; echo "bench exit $?"
cut -c1-330 gpurun_out/r35_bench.json
