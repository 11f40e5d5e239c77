#!/bin/bash
# round 2, call 16 (1 GPU): fc1 with the activation in its own epilogue (fast_gemm_act.cu, M = 64 tcgen05 tiles over the
# whole K, X through a ring) against the split-K + reduce/activation pair (varF0 = -DPTTS_FC1_ACT=0), same box;
# varP2 = -DPTTS_ATTN_PRE2=1 (self-attention requests two K batches before the dependency wait);
# elect.sync in the non-lean GEMM (heads GEMM, prefill, T5); per-layer table of the DAC decoder; kernel table of T5 + prefill
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fast_kernels.py tests/test_gpu_parity.py tests/test_gpu_baseline_configs.py tests/test_gpu_t5.py -m gpu -x -q > gpurun_out/r2_16_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_16_pytest.log
tail -15 gpurun_out/r2_16_pytest.log
line() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ms/step", round(d["ms_per_decode_step"],4), "value", round(d["value"]), "frac", round(d["roofline"]["frac"],3),
          "launches/step", d.get("gpu_launches"), "check", d["check"].get("ok"), d["check"].get("codes_sha16"), d["check"].get("simple_kernels_logit_rel_err"))
except Exception as e:
    print("   no line:", e)
PY
}
cp parler_tts_b200/libparler_b200.so /tmp/default.so
for v in default varF0 varP2 default varF0 varP2; do
  if [ $v = default ]; then cp /tmp/default.so parler_tts_b200/libparler_b200.so; else cp parler_tts_b200/libparler_b200_$v.so parler_tts_b200/libparler_b200.so; fi
  timeout 900 python bench.py --skip-cpu --skip-e2e > gpurun_out/r2_16_bench_$v.json 2> gpurun_out/r2_16_bench_$v.err; echo "$v rc=$?"; line gpurun_out/r2_16_bench_$v.json
done
cp /tmp/default.so parler_tts_b200/libparler_b200.so
timeout 900 python bench.py --skip-cpu --skip-e2e --kv-dtype fp8 > gpurun_out/r2_16_bench_fp8.json 2> gpurun_out/r2_16_bench_fp8.err; echo "fp8 rc=$?"; line gpurun_out/r2_16_bench_fp8.json
timeout 600 python profiles/scripts/dac_layers.py > gpurun_out/r2_16_dac_layers.txt 2>&1; tail -32 gpurun_out/r2_16_dac_layers.txt
timeout 600 python profiles/scripts/prefill_profile.py > gpurun_out/r2_16_prefill_profile.txt 2>&1; grep -v "^-" gpurun_out/r2_16_prefill_profile.txt | cut -c1-72,150-260 | head -44
