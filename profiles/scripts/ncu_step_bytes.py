"""Summarises the ncu CSV of profiles/scripts/ncu_step.py (one eager decode step): DRAM bytes and time per kernel class
and for the whole step, against the algorithmic bytes of SURVEY.md section 8d.  Writes profiles/r02_step_dram_bytes.json
(read by bench.py for roofline.traffic) and prints a table.
    python profiles/scripts/ncu_step_bytes.py gpurun_out/step_bytes.csv bf16 > profiles/r02_step_dram_bytes_bf16.txt
"""
import collections
import csv
import json
import os
import re
import sys

path, kv = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "bf16")
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if not ln.startswith("==")]
for r in csv.DictReader(lines):
    rows.append(r)
per = collections.defaultdict(lambda: collections.defaultdict(float))
launch = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"^void |ptts::|\(anonymous namespace\)::", "", name)
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    m = r["Metric Name"]
    if m.startswith("dram__bytes") or m.startswith("lts__t_bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    if m == "gpu__time_duration.sum":
        v *= {"nsecond": 1e-3, "usecond": 1, "msecond": 1e3, "ns": 1e-3, "us": 1, "ms": 1e3}.get(unit, 1)
    per[name][m] += v
    launch[r["ID"]] = name
counts = collections.Counter(launch.values())
tot = collections.defaultdict(float)
print(f"{'kernel':70s} {'launches':>8s} {'time us':>9s} {'DRAM rd MB':>11s} {'DRAM wr MB':>11s} {'L2 MB':>9s}")
for name, m in sorted(per.items(), key=lambda kv_: -kv_[1]["gpu__time_duration.sum"]):
    print(f"{name[:70]:70s} {counts[name]:8d} {m['gpu__time_duration.sum']:9.1f} {m['dram__bytes_read.sum'] / 1e6:11.2f} "
          f"{m['dram__bytes_write.sum'] / 1e6:11.2f} {m['lts__t_bytes.sum'] / 1e6:9.1f}")
    for k, v in m.items():
        tot[k] += v
B, S, L, H, Kc, ctx = 64, 48, 24, 1024, 9, 466
W = 362498048
es_kv = 1 if kv == "fp8" else 2
alg = W * 2 + B * (2 * L * H * es_kv * ctx + 2 * L * H * 2 * S) + B * 2 * L * H * es_kv + B * Kc * H * 4
dram = tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]
print(f"\nwhole step: {len(launch)} launches, {tot['gpu__time_duration.sum']:.1f} us serialised under ncu, DRAM read "
      f"{tot['dram__bytes_read.sum'] / 1e6:.1f} MB + write {tot['dram__bytes_write.sum'] / 1e6:.1f} MB = {dram / 1e6:.1f} MB; "
      f"algorithmic {alg / 1e6:.1f} MB at context {ctx} -> traffic / algorithmic = {dram / alg:.3f}")
out = {"kv_dtype": kv, "context": ctx, "launches": len(launch), "dram_bytes_per_step": dram,
       "dram_read": tot["dram__bytes_read.sum"], "dram_write": tot["dram__bytes_write.sum"],
       "l2_bytes": tot["lts__t_bytes.sum"], "algorithmic_bytes": alg, "ratio": dram / alg,
       "source": "ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,... "
                 "python profiles/scripts/ncu_step.py " + kv}
if kv == "bf16":
    json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                     "r02_step_dram_bytes.json"), "w"), indent=1)
