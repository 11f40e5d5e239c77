# bench lines for the other BASELINE configs on the current build: configs[4] (B=16, 2600 frames, incl. DAC through e2e),
# per-GPU batch 128 (configs[2] at 4 GPUs), end-to-end phase split, and the default bench line of record
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 1 --skip-cpu --batch 16 --frames 2600 --e2e-steps 1 > gpurun_out/r8_long.json 2> gpurun_out/r8_long.err; echo "long exit $?"
timeout 400 python bench.py --steps 1 --warmup 1 --skip-e2e --skip-cpu --batch 128 > gpurun_out/r8_b128.json 2> gpurun_out/r8_b128.err; echo "b128 exit $?"
timeout 400 python profiles/scripts/e2e_phases.py > gpurun_out/r8_e2e_phases.txt 2>&1; echo "phases exit $?"
tail -3 gpurun_out/r8_e2e_phases.txt
timeout 600 python bench.py > gpurun_out/r8_bench.json 2> gpurun_out/r8_bench.err; echo "bench exit $?"
cat gpurun_out/r8_bench.json
for f in long b128; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r8_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["value"]), "frames/s", round(d["ms_per_decode_step"], 4), "ms/step", "frac", round(d["roofline"]["whole_step"]["frac"], 3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/r8_$f.err").read()[-1500:])
PY
done
