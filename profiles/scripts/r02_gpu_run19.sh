#!/bin/bash
# round 2, call 19 (1 GPU): prefill attention with staged K / V (fast_attn_prefill.cu), staged T5 attention, heads
# partials with an unrolled slice fold (default) against varH0 (STORE GEMM + logits matrix); full GPU suite
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_19_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_19_pytest.log
tail -12 gpurun_out/r2_19_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "e2e", d["e2e"] and round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],3),
          "check", d["check"].get("ok"), d["check"].get("codes_sha16"), d["check"].get("eager_replay_token_match"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varH0 default varH0; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e > gpurun_out/r2_19_bench_$v.json 2> gpurun_out/r2_19_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_19_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
timeout 600 python profiles/scripts/e2e_phases2.py > gpurun_out/r2_19_e2e_phases2.txt 2>&1; tail -12 gpurun_out/r2_19_e2e_phases2.txt
timeout 600 python profiles/scripts/prefill_profile.py > gpurun_out/r2_19_prefill_profile.txt 2>&1; grep -v "^-" gpurun_out/r2_19_prefill_profile.txt | cut -c1-72,150-260 | grep "ptts\|====" | head -24
