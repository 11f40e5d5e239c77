"""Per-kernel summary of an `ncu --set full` raw-page CSV (ncu -i x.ncu-rep --page raw --csv):
    python profiles/scripts/ncu_full_summary.py gpurun_out/r2_3_step_full.raw.csv > profiles/r02_ncu_full_kernels.txt
One line per distinct (kernel, grid) with the metrics B200_PROFILING.md names: duration, DRAM bytes and throughput, tensor
pipe activity, bf16 tensor ops per second against the pipe's sustained peak, warps active, registers, waves."""
import collections
import csv
import re
import sys

rows = list(csv.DictReader(open(sys.argv[1])))
units = next((r for r in rows if not r["ID"].strip()), {})  # the raw page's second line holds the units
rows = [r for r in rows if r.get("Kernel Name") and r["ID"].strip()]
COLS = [("gpu__time_duration.sum", "time", 1), ("dram__bytes_read.sum", "DRAM rd", 1), ("dram__bytes_write.sum", "DRAM wr", 1),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", 1),
        ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe %", 1),
        ("sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.sum.pct_of_peak_sustained_elapsed", "bf16 MMA % of peak", 1),
        ("sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.sum.per_second", "bf16 MMA op/s", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %", 1),
        ("launch__registers_per_thread", "regs", 1), ("launch__grid_size", "grid", 1), ("launch__block_size", "block", 1),
        ("launch__waves_per_multiprocessor", "waves/SM", 1)]
data = rows
groups = collections.OrderedDict()
for r in data:
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"^void |ptts::|\(anonymous namespace\)::", "", name)
    key = (name, r.get("launch__grid_size", ""))
    groups.setdefault(key, []).append(r)


def num(v):
    try:
        return float(v.replace(",", ""))
    except Exception:
        return float("nan")


print("ncu --set full, per distinct (kernel, grid): mean over the launches captured; units as ncu reports them")
print("units:", {lbl: units.get(k, "") for k, lbl, _ in COLS if units.get(k, "")})
for (name, grid), rs in groups.items():
    print(f"\n{name}  x{len(rs)}")
    parts = []
    for k, lbl, _ in COLS:
        vals = [num(r.get(k, "nan")) for r in rs]
        vals = [v for v in vals if v == v]
        if vals:
            m = sum(vals) / len(vals)
            parts.append(f"{lbl} {m:.4g}")
    print("   " + "; ".join(parts))
