#!/bin/bash
# round 2, call 36 (1 GPU): ncu --set full of DAC conv launches at a realistic size (B=8 x 860 frames): block 2 (384
# channels) k7 + k1, block 4 (96 channels) k7 + k1 - which pipe bounds them?
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python profiles/scripts/dac_ncu.py > gpurun_out/r2_36_plain.log 2>&1; tail -1 gpurun_out/r2_36_plain.log
for sel in "9 2 b2" "23 2 b4"; do set -- $sel
  timeout 900 ncu --profile-from-start off --clock-control none --set full -k regex:dac_conv_ws --launch-skip $1 --launch-count $2 \
    -o gpurun_out/r2_36_dac_$3 python profiles/scripts/dac_ncu.py > gpurun_out/r2_36_ncu_$3.log 2>&1; echo "ncu $3 rc=$?"
  ncu -i gpurun_out/r2_36_dac_$3.ncu-rep --page raw --csv > gpurun_out/r2_36_dac_$3.raw.csv 2>/dev/null
  ncu -i gpurun_out/r2_36_dac_$3.ncu-rep --page details > gpurun_out/r2_36_dac_$3.details.txt 2>/dev/null
  rm -f gpurun_out/r2_36_dac_$3.ncu-rep
done
ls -la gpurun_out | grep r2_36
