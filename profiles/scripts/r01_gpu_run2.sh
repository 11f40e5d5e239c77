# A/B of the next-layer K prefetch placement, top-k sampling epilogue, per-GPU batch 256 (configs[2] at 2 GPUs) and the
# scaled decoder (configs[3]); bench lines only (no e2e / CPU legs)
set -x
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --skip-e2e --skip-cpu"
PTTS_L2PF_LATE=0 timeout 300 $B > gpurun_out/r2_pf_early.json 2> gpurun_out/r2_pf_early.err; echo "exit $?"
PTTS_L2PF_LATE=1 timeout 300 $B > gpurun_out/r2_pf_late.json 2> gpurun_out/r2_pf_late.err; echo "exit $?"
timeout 300 $B --do-sample > gpurun_out/r2_sample.json 2> gpurun_out/r2_sample.err; echo "exit $?"
timeout 400 python bench.py --steps 1 --warmup 1 --skip-e2e --skip-cpu --batch 256 > gpurun_out/r2_b256.json 2> gpurun_out/r2_b256.err; echo "exit $?"
timeout 500 python bench.py --steps 1 --warmup 1 --skip-cpu --arch scaled --batch 32 --e2e-steps 1 > gpurun_out/r2_scaled.json 2> gpurun_out/r2_scaled.err; echo "exit $?"
for f in pf_early pf_late sample b256 scaled; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["value"]), "frames/s", round(d["ms_per_decode_step"], 4), "ms/step", "whole-step frac", round(d["roofline"]["whole_step"]["frac"], 3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/r2_$f.err").read()[-1500:])
PY
done
