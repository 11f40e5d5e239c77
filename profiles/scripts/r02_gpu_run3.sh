#!/bin/bash
# round 2, call 3: full GPU test suite, then the ncu passes (step-level DRAM bytes for both cache dtypes; --set full of
# the step's kernels + the DAC conv kernels on a 2-layer model of the same widths)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2_3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_3_pytest.log
tail -4 gpurun_out/r2_3_pytest.log
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum
for kv in bf16 fp8; do
  timeout 600 python profiles/scripts/ncu_step.py $kv > gpurun_out/r2_3_plain_$kv.log 2>&1 &&
  timeout 1500 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/r2_3_step_bytes_$kv.csv \
    --metrics $M python profiles/scripts/ncu_step.py $kv > gpurun_out/r2_3_ncu_$kv.log 2>&1
  echo "ncu step bytes $kv rc=$?"
done
timeout 600 python profiles/scripts/ncu_step.py bf16 2 dac > gpurun_out/r2_3_plain_full.log 2>&1 &&
timeout 1800 ncu --profile-from-start off --clock-control none --set full -o gpurun_out/r2_3_step_full \
  python profiles/scripts/ncu_step.py bf16 2 dac > gpurun_out/r2_3_ncu_full.log 2>&1
echo "ncu set full rc=$?"
timeout 600 python profiles/scripts/ncu_step.py fp8 2 > gpurun_out/r2_3_plain_full8.log 2>&1 &&
timeout 1800 ncu --profile-from-start off --clock-control none --set full -k regex:attn_fp8 -o gpurun_out/r2_3_attn_fp8_full \
  python profiles/scripts/ncu_step.py fp8 2 > gpurun_out/r2_3_ncu_full8.log 2>&1
echo "ncu set full fp8 rc=$?"
for f in r2_3_step_full r2_3_attn_fp8_full; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null
  ls -la gpurun_out/$f.ncu-rep
  # keep the report only if it fits the 64 MiB return channel with room to spare
  [ "$(stat -c %s gpurun_out/$f.ncu-rep)" -gt 30000000 ] && rm -f gpurun_out/$f.ncu-rep
done
ls -la gpurun_out | tail -20
