set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1_pytest.log
tail -5 gpurun_out/r1_pytest.log
timeout 600 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench exit $?"
cat gpurun_out/r1_bench.json
timeout 600 python profiles/scripts/microbatch.py > gpurun_out/r1_microbatch.txt 2>&1; echo "mb exit $?"
cat gpurun_out/r1_microbatch.txt | tail -12
timeout 300 python profiles/scripts/attn_ncu.py > gpurun_out/r1_attn_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 24 -c 3 -o gpurun_out/r1_attn_full python profiles/scripts/attn_ncu.py > gpurun_out/r1_attn_ncu.log 2>&1
echo "ncu full exit $?"; cat gpurun_out/r1_attn_plain.log
timeout 300 python bench.py --steps 1 --warmup 1 --frames 100 --skip-e2e --skip-cpu --no-graph > gpurun_out/r1_short_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 40000 -c 900 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 1 --warmup 1 --frames 100 --skip-e2e --skip-cpu --no-graph > gpurun_out/r1_short_ncu.log 2>&1
echo "ncu list exit $?"
