# bench lines of record (v8): default line, reference arm, sampling / long-form / scaled-decoder lines, ncu launch list of a
# mid-context stretch of the eager step sequence, ncu --set full captures of the cross-attention block, the GEMM and
# the flash-decode kernel
set -x
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r19_bench.json 2> gpurun_out/r19_bench.err; echo "bench exit $?"
cat gpurun_out/r19_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r19_reference.json 2> gpurun_out/r19_reference.err; echo "reference exit $?"
cat gpurun_out/r19_reference.json
timeout 400 python bench.py --steps 2 --warmup 1 --skip-cpu --do-sample --e2e-steps 1 > gpurun_out/r19_sample.json 2> gpurun_out/r19_sample.err; echo "sample exit $?"
timeout 600 python bench.py --steps 1 --warmup 1 --skip-cpu --batch 16 --frames 2600 --e2e-steps 1 > gpurun_out/r19_long.json 2> gpurun_out/r19_long.err; echo "long exit $?"
timeout 600 python bench.py --steps 1 --warmup 1 --skip-cpu --skip-e2e --arch scaled --batch 32 > gpurun_out/r19_scaled.json 2> gpurun_out/r19_scaled.err; echo "scaled exit $?"
for f in sample long scaled; do python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r19_$f.json").read().strip().splitlines()[-1])
    print("$f", round(d["value"]), "frames/s", round(d["ms_per_decode_step"], 4), "ms/step", "frac", round(d["roofline"]["whole_step"]["frac"], 3), "e2e", d.get("e2e") and round(d["e2e"]["value"]))
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/r19_$f.err").read()[-1500:])
PY
done
# launch list: eager (no graph) so that ncu sees every kernel; skip into the middle of the timed generation (context ~470)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 290000 -c 900 --csv --log-file gpurun_out/r19_launches.csv python bench.py --steps 1 --warmup 1 --skip-e2e --skip-cpu --no-graph > gpurun_out/r19_launches_run.log 2>&1; echo "ncu list exit $?"
tail -3 gpurun_out/r19_launches.csv | cut -c1-200
timeout 600 ncu --set full --clock-control none --import-source on -k regex:xblock -s 2000 -c 3 -o gpurun_out/r19_xblock_full python bench.py --steps 1 --warmup 0 --frames 200 --skip-e2e --skip-cpu --no-graph > gpurun_out/r19_xblock_ncu.log 2>&1; echo "ncu xblock exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_gemm -s 8000 -c 4 -o gpurun_out/r19_gemm_full python bench.py --steps 1 --warmup 0 --frames 200 --skip-e2e --skip-cpu --no-graph > gpurun_out/r19_gemm_ncu.log 2>&1; echo "ncu gemm exit $?"
timeout 300 python profiles/scripts/attn_ncu.py > gpurun_out/r19_attn_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 24 -c 3 -o gpurun_out/r19_attn_full python profiles/scripts/attn_ncu.py > gpurun_out/r19_attn_ncu.log 2>&1; echo "ncu attn exit $?"
cat gpurun_out/r19_attn_plain.log
ls -la gpurun_out/*.ncu-rep
