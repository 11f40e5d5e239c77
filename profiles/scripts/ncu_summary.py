"""Condenses an `ncu --set full` report into the small CSV kept under profiles/ (one column per captured launch).
  python profiles/scripts/ncu_summary.py gpurun_out/x.ncu-rep "header line" > profiles/x.csv"""
import csv
import io
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__cluster_size", "launch__shared_mem_per_block_dynamic",
        "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg"]
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
name_col = hdr.index("Kernel Name")
print("# " + (sys.argv[2] if len(sys.argv) > 2 else rep))
print("# kernels: " + " | ".join(sorted({r[name_col][:110] for r in data})))
print("metric,unit," + ",".join(f"launch{i}" for i in range(len(data))))
for k in KEEP:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k},{units[i]}," + ",".join(r[i] for r in data))
