"""Registers / local-memory frame / static shared memory per hot kernel of the built library:
    cuobjdump -res-usage parler_tts_b200/libparler_b200.so | python profiles/scripts/res_usage_summary.py > profiles/r02_res_usage_v21.txt
"""
import re
import subprocess
import sys

txt = sys.stdin.read()
rows = [(m.group(1), m.group(2).strip()) for m in re.finditer(r"Function (\S+):\n\s*(.*)", txt)]
names = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.split("\n")
print("Resource usage per hot kernel of parler_tts_b200/libparler_b200.so: cuobjdump -res-usage.\n"
      "REG = registers per thread, STACK > 0 = local-memory frame (spills or local arrays), SHARED = static shared memory.\n")
hot = ("attn_decode_kernel", "umma_partial_lean", "umma_act", "xblock_kernel", "ln_reduce", "step_epilogue", "dac_conv_ws",
       "gemm_big", "attn_fp8", "attn_prefill", "t5_attention_staged", "dac_conv_out", "attn_noalloc")
for (mangled, res), name in zip(rows, names):
    if any(h in name for h in hot):
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("void ", "").replace("ptts::", "")
        print(f"{short[:78]:78s} {res}")
