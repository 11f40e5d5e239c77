"""Marginal in-graph cost of each kernel class of the decode step (timing only: a skipped class leaves garbage):
diagnostic library (make DIAG=1) with PTTS_SKIP=<mask>, one subprocess per mask, B=64, replays at context ~466.
    python profiles/scripts/ablate2.py            (parent: loops over the masks)
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
MASKS = [(0, "nothing skipped"), (1, "LayerNorms (48 + 1)"), (2, "QKV GEMM"), (4, "self-attention"),
         (8, "out_proj + fc2 GEMMs"), (256, "cross-attention block"), (64, "fc1 + GELU"), (2 | 8 | 64, "all GEMMs"),
         (1 | 2 | 8 | 64 | 256, "everything but self-attention"), (4 | 256, "both attentions")]

if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch

    sys.path.insert(0, ROOT)
    import bench
    from parler_tts_b200.engine import DecoderEngine, GenState

    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    B, S, P, K, M = 64, 48, 32, 9, 869
    model = bench.build_models(dev, torch.bfloat16)
    w = model.decoder.packed_weights()
    eng = DecoderEngine(w)
    st = GenState(w, B, S, P, 1, M, suppress_from=1024)
    enc = (torch.randn(B, S, 1024, device=dev) * 0.3).to(torch.bfloat16)
    prompt = torch.randint(2, 32100, (B, P), device=dev)
    st.reset(torch.full((B * K, 1), 1025, dtype=torch.int64, device=dev), seed=0)
    eng.cross_kv(st, enc)
    eng.prefill(st, model.prompt_table(), prompt)
    eng.run_loop(st, use_graph=True, poll_every=64, max_steps=420)  # context 33 + 421 = 454
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(24):  # contexts 455 .. 478 (mean 466)
        st.graph.replay()
    b.record()
    torch.cuda.synchronize()
    print("MS_PER_STEP %.4f %d" % (a.elapsed_time(b) / 24, st.graph_launches))
    sys.exit(0)

base = None
for mask, name in MASKS:
    env = dict(os.environ, PTTS_LIBRARY=os.path.join(ROOT, "parler_tts_b200", "libparler_b200_diag.so"), PTTS_SKIP=str(mask))
    out = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True).stdout
    line = [ln for ln in out.splitlines() if ln.startswith("MS_PER_STEP")]
    if not line:
        print(f"skip {mask:4d} {name}: failed")
        continue
    ms, launches = float(line[0].split()[1]), int(line[0].split()[2])
    if base is None:
        base = ms
    print(f"skip {mask:4d}  {name:34s} {ms:7.4f} ms per step  ({launches:3d} launches)   marginal cost of the skipped class {base - ms:+.4f} ms")
