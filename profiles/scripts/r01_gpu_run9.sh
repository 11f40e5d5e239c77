set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_generate.py -m gpu -x -q -k "encode or audio_prompt" > gpurun_out/r9_pytest_new.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r9_pytest_new.log | cut -c1-250
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r9_pytest_all.log 2>&1; echo "pytest-all exit $?"
tail -8 gpurun_out/r9_pytest_all.log | cut -c1-250
