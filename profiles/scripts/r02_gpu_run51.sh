#!/bin/bash
# round 2, call 51 (1 GPU): ncu launch list (serialised durations) of T5 encoder + cross K/V + prefill
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2_51_prefill_launches.csv python profiles/scripts/prefill_ncu.py > gpurun_out/r2_51_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r2_51_ncu.log
python - <<'PY'
import csv, collections, re
rows = [r for r in csv.reader(open("gpurun_out/r2_51_prefill_launches.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("ptts::", "").replace("(anonymous namespace)::", "").replace("void ", "")
    grid = r[7] if len(r) > 7 else ""
    val = float(r[-1].replace(",", ""))
    unit = r[-2]
    us = val / 1000.0 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1000.0)
    k = (name[:70], grid)
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"launches {len(rows)}, sum of serialised durations {tot / 1000:.2f} ms")
for (name, grid), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"  {name:70s} grid {grid:16s} n={n:4d}  {us / n:8.1f} us each  {us / 1000:7.2f} ms")
PY
