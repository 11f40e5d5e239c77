#!/bin/bash
# round 2, call 61 (1 GPU): ncu --set full of the 128 x 128 tile GEMM (prefill shapes) and of the 32-row fc1 kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_big_kernel -s 60 -c 8 -o gpurun_out/r2_61_gemm_big -f python profiles/scripts/prefill_ncu.py > gpurun_out/r2_61_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r2_61_ncu.log
ncu -i gpurun_out/r2_61_gemm_big.ncu-rep --page raw --csv > gpurun_out/r2_61_gemm_big_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/r2_61_gemm_big_raw.csv")))
hdr = rows[0]
want = ["Kernel Name", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_uniform.sum", "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_cycles_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active", "launch__registers_per_thread", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
idx = [(w, hdr.index(w)) for w in want if w in hdr]
tens = [h for h in hdr if "tensor" in h][:12]
print("tensor-ish metrics:", tens)
print("units:", [(w, rows[1][i]) for w, i in idx])
for r in rows[2:]:
    print({w: r[i][:40] for w, i in idx})
PY
