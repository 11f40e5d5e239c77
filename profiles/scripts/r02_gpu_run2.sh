#!/bin/bash
# round 2, call 2: full GPU test suite (no -x), L2 prefetch A/B for both cache dtypes
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s --maxfail=40 > gpurun_out/r2_2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_2_pytest.log
tail -5 gpurun_out/r2_2_pytest.log
for kv in fp8 bf16; do
  for pf in 0 50 100; do
    timeout 600 python bench.py --steps 2 --warmup 2 --kv-dtype $kv --kv-prefetch $pf --skip-cpu --skip-e2e \
      > gpurun_out/r2_2_bench_${kv}_pf${pf}.json 2> gpurun_out/r2_2_bench_${kv}_pf${pf}.err
    echo "bench $kv pf=$pf rc=$?"
    python - <<PY
import json
d=json.loads(open("gpurun_out/r2_2_bench_${kv}_pf${pf}.json").read().strip().splitlines()[-1])
print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "check", d["check"].get("ok"))
PY
  done
done
