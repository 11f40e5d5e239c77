#!/bin/bash
# round 2, call 20 (1 GPU): evidence refresh on the v14 library (195 launches per step): streaming tests, step-level DRAM
# bytes for both cache dtypes, ncu --set full of the step's kernels + DAC convs on a 2-layer model of the same widths,
# launch list (gpu__time_duration) of one eager step
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_streaming.py -m gpu -x -q > gpurun_out/r2_20_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_20_pytest.log
tail -3 gpurun_out/r2_20_pytest.log
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum
for kv in bf16 fp8; do
  timeout 600 python profiles/scripts/ncu_step.py $kv > gpurun_out/r2_20_plain_$kv.log 2>&1 &&
  timeout 1500 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/r2_20_step_bytes_$kv.csv \
    --metrics $M python profiles/scripts/ncu_step.py $kv > gpurun_out/r2_20_ncu_$kv.log 2>&1
  echo "ncu step bytes $kv rc=$?"
done
timeout 600 python profiles/scripts/ncu_step.py bf16 2 dac > gpurun_out/r2_20_plain_full.log 2>&1 &&
timeout 1800 ncu --profile-from-start off --clock-control none --set full --import-source on -o gpurun_out/r2_20_step_full \
  python profiles/scripts/ncu_step.py bf16 2 dac > gpurun_out/r2_20_ncu_full.log 2>&1
echo "ncu set full rc=$?"
ncu -i gpurun_out/r2_20_step_full.ncu-rep --page raw --csv > gpurun_out/r2_20_step_full.raw.csv 2>/dev/null
ls -la gpurun_out/r2_20_step_full.ncu-rep
[ "$(stat -c %s gpurun_out/r2_20_step_full.ncu-rep)" -gt 30000000 ] && rm -f gpurun_out/r2_20_step_full.ncu-rep
ls -la gpurun_out | tail -12
