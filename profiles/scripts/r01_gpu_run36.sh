# chunked (streaming) DAC decode equals the one-shot decode
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_generate.py -m gpu -x -q -k "chunked or dac_44k" > gpurun_out/r36_pytest.log 2>&1; echo "pytest exit $?"
tail -12 gpurun_out/r36_pytest.log | cut -c1-300
