"""Codec decode beside the token loop without a streamer (model.overlap_decode_chunk_frames / overlap_decode_ctas): wall
time of generate() from host ids to host audio at the benched size (Mini v0.1, B=64, 860 frames, bf16), against the
decode-after-the-loop call on the same box.
    python profiles/scripts/overlap_time.py [chunk:ctas ...]      (ctas 0 = one per SM)"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200.engine import DecoderEngine  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = int(os.environ.get("OVL_B", 64)), 48, 32, 860
combos = [tuple(int(v) for v in a.split(":")) for a in sys.argv[1:]] or [(128, 0), (128, 48), (128, 64), (256, 48), (256, 64)]
model = bench.build_models(dev, torch.bfloat16)
model.generation_config.max_length = N + 9
g = torch.Generator().manual_seed(0)
ids = torch.randint(2, 32100, (B, S), generator=g).pin_memory()
pids = torch.randint(2, 32100, (B, P), generator=g).pin_memory()
procs = [ptts.SuppressTokensAtOrAbove(1024)]


def run():
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    audio = model.generate(input_ids=ids.to(dev, non_blocking=True), prompt_input_ids=pids.to(dev, non_blocking=True),
                           do_sample=False, logits_processor=procs, seed=0)
    host = audio.cpu()
    return (time.perf_counter() - t0) * 1e3, host


def best(n=3):
    out = [run() for _ in range(n)]
    return min(o[0] for o in out), out[-1][1]


def timeline(cf, ctas):
    """one traced call: when each chunk decode ran on the second stream, when the loop ended (ms after the prefill)"""
    st = ptts.streaming.SilentChunks(cf, ctas or None)
    st.trace = True
    torch.cuda.synchronize()
    model.generate(input_ids=ids.to(dev), prompt_input_ids=pids.to(dev), do_sample=False, logits_processor=procs, seed=0,
                   streamer=st)
    tl = st.timeline
    print(f"  timeline chunk {cf} / {ctas} CTAs: loop ended at {tl['loop_end_ms']:7.1f} ms, call's device work at {tl['end_ms']:7.1f} ms")
    for f0, f1, steps, a, b in tl["chunks"]:
        print(f"    frames [{f0:4d}, {f1:4d}) queued at step {steps:4d}: decode ran {a:7.1f} -> {b:7.1f} ms ({b - a:6.1f} ms)")


for prio in [int(v) for v in os.environ.get('OVL_PRIO', '0,-1').split(',')]:
    DecoderEngine.graph_priority = prio
    model.decoder._state_cache = None  # capture again with the new priority
    model.overlap_decode_chunk_frames, model.overlap_decode_ctas = 0, None
    run(); run()
    base, want = best()
    print(f"graph priority {prio:2d}: decode after the loop {base:8.1f} ms per call, {B * N / base * 1e3:9.0f} frames/s")
    for cf, ctas in combos:
        model.overlap_decode_chunk_frames, model.overlap_decode_ctas = cf, (ctas or None)
        run()
        ms, host = best()
        err = float((host - want).abs().max())
        print(f"  chunk {cf:4d} frames, {ctas:3d} CTAs: {ms:8.1f} ms ({ms / base:5.3f} x), {B * N / ms * 1e3:9.0f} frames/s, "
              f"max |diff| {err:.2e}")
        if os.environ.get("OVL_TIMELINE"):
            timeline(cf, ctas)
