#!/bin/bash
# round 2, call 69 (1 GPU): evidence refresh on the v21 library - step-level DRAM bytes of one eager step (bf16 cache),
# ncu --set full of the step's kernels + DAC convs on a 2-layer model of the same widths
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum
timeout 300 python profiles/scripts/ncu_step.py bf16 > gpurun_out/r2_69_plain_bf16.log 2>&1 &&
timeout 600 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/r2_69_step_bytes_bf16.csv \
  --metrics $M python profiles/scripts/ncu_step.py bf16 > gpurun_out/r2_69_ncu_bf16.log 2>&1
echo "ncu step bytes bf16 rc=$?"
timeout 300 python profiles/scripts/ncu_step.py bf16 2 dac > gpurun_out/r2_69_plain_full.log 2>&1 &&
timeout 900 ncu --profile-from-start off --clock-control none --set full --import-source on -o gpurun_out/r2_69_step_full \
  python profiles/scripts/ncu_step.py bf16 2 dac > gpurun_out/r2_69_ncu_full.log 2>&1
echo "ncu set full rc=$?"
ncu -i gpurun_out/r2_69_step_full.ncu-rep --page raw --csv > gpurun_out/r2_69_step_full.raw.csv 2>/dev/null
ls -la gpurun_out/r2_69_step_full.ncu-rep
[ "$(stat -c %s gpurun_out/r2_69_step_full.ncu-rep)" -gt 30000000 ] && rm -f gpurun_out/r2_69_step_full.ncu-rep
ls -la gpurun_out | tail -8
