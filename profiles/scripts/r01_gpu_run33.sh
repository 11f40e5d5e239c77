# state of record at the end of round 1: whole GPU suite, smoke(), default bench line, reference arm, timeline, phases
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r33_pytest_all.log 2>&1; echo "pytest-all exit $?"
tail -4 gpurun_out/r33_pytest_all.log | cut -c1-300
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r33_smoke.log 2>&1; echo "smoke exit $?"
tail -2 gpurun_out/r33_smoke.log
timeout 900 python bench.py > gpurun_out/r33_bench.json 2> gpurun_out/r33_bench.err; echo "bench exit $?"
cat gpurun_out/r33_bench.json | cut -c1-300
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r33_reference.json 2> gpurun_out/r33_reference.err; echo "reference exit $?"
cut -c1-200 gpurun_out/r33_reference.json
