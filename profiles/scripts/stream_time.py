"""Streaming generate() at the benched size (Mini v0.1, B=64, 860 frames, bf16): wall time until the first audio chunk
is in pinned host memory, total time of the call, against the one-shot generate() of the same inputs.
    python profiles/scripts/stream_time.py [chunk_frames ...]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import parler_tts_b200 as ptts  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, 860
chunks = [int(a) for a in sys.argv[1:]] or [64, 128, 256]
model = bench.build_models(dev, torch.bfloat16)
model.generation_config.max_length = N + 9
g = torch.Generator().manual_seed(0)
ids = torch.randint(2, 32100, (B, S), generator=g).pin_memory()
pids = torch.randint(2, 32100, (B, P), generator=g).pin_memory()
procs = [ptts.SuppressTokensAtOrAbove(1024)]


def run(streamer=None):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    audio = model.generate(input_ids=ids.to(dev, non_blocking=True), prompt_input_ids=pids.to(dev, non_blocking=True),
                           do_sample=False, logits_processor=procs, seed=0, streamer=streamer)
    host = audio.cpu()
    return (time.perf_counter() - t0) * 1e3, host, t0


for _ in range(2):
    base_ms, want, _ = run()
base = min(run()[0] for _ in range(3))
print(f"one-shot generate(): {base:8.1f} ms per call (host ids -> host audio), {B * N / base * 1e3:9.0f} frames/s")
for cf in chunks:
    for rep in range(3):
        first = []
        st = ptts.AudioChunkStreamer(chunk_frames=cf, on_chunk=lambda f, a: first.append(time.perf_counter()))
        ms, host, t0 = run(st)
        if rep == 0:
            continue  # warm-up (side-stream allocations)
        err = float((host - want).abs().max())
        print(f"chunk {cf:4d} frames: first audio after {(first[0] - t0) * 1e3:7.1f} ms, {len(first):3d} chunks, call "
              f"{ms:8.1f} ms ({ms / base:5.3f} x one-shot), {B * N / ms * 1e3:9.0f} frames/s, max |diff| vs one-shot {err:.2e}")
