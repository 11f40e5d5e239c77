# new defaults (early release, no own-V prefetch) + sweep of the next-layer prefetch share (100 = all K rows, 200 = K and V)
set -x
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu"
run() { name=$1; shift; env "$@" timeout 300 $B > gpurun_out/r5_$name.json 2> gpurun_out/r5_$name.err; echo "$name exit $?"; }
run base PTTS_X=0
run pct50 PTTS_L2PF_PCT=50
run pct130 PTTS_L2PF_PCT=130
run pct160 PTTS_L2PF_PCT=160
run pct200 PTTS_L2PF_PCT=200
run nopf PTTS_L2PF=0
timeout 300 python profiles/scripts/step_timeline.py > gpurun_out/r5_timeline.txt 2>&1; echo "timeline exit $?"
cat gpurun_out/r5_timeline.txt | tail -22
for f in base pct50 pct130 pct160 pct200 nopf; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r5_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["value"]), "frames/s", round(d["ms_per_decode_step"], 4), "ms/step")
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/r5_$f.err").read()[-1500:])
PY
done
