set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_t5.py tests/test_gpu_fast_kernels.py -m gpu -x -q -k "t5 or scaled_and_mini" > gpurun_out/r3_pytest_new.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r3_pytest_new.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r3_pytest_all.log 2>&1; echo "pytest-all exit $?"
tail -8 gpurun_out/r3_pytest_all.log
timeout 600 python bench.py --steps 2 --warmup 1 --skip-cpu > gpurun_out/r3_bench.json 2> gpurun_out/r3_bench.err; echo "bench exit $?"
cat gpurun_out/r3_bench.json; tail -5 gpurun_out/r3_bench.err
