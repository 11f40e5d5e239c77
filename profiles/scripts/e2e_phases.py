"""Where does the end-to-end generate() time go?  Mini v0.1, B=64, 860 frames, bf16 decoder: CUDA-event timing of
(a) text encoder + projections + prefill + decode loop (generate(return_codes=True)) and (b) de-delay'd codes -> DAC."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import parler_tts_b200 as ptts  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, int(sys.argv[1]) if len(sys.argv) > 1 else 860
model = bench.build_models(dev, torch.bfloat16)
model.generation_config.max_length = N + 9
g = torch.Generator().manual_seed(0)
ids = torch.randint(2, 32100, (B, S), generator=g).pin_memory()
pids = torch.randint(2, 32100, (B, P), generator=g).pin_memory()
procs = [ptts.SuppressTokensAtOrAbove(1024)]


def ev():
    return torch.cuda.Event(enable_timing=True)


for it in range(2):
    e = [ev() for _ in range(4)]
    e[0].record()
    codes = model.generate(input_ids=ids.to(dev, non_blocking=True), prompt_input_ids=pids.to(dev, non_blocking=True),
                           do_sample=False, logits_processor=procs, seed=0, return_codes=True)
    e[1].record()
    gc = model.generation_config
    audio = model.audio_encoder.engine(model.decoder.impl).decode_generated(codes, 1025, 1024, 1024)
    e[2].record()
    host = audio.cpu()
    e[3].record()
    torch.cuda.synchronize()
    print(f"run {it}: codes {tuple(codes.shape)} generate(return_codes) {e[0].elapsed_time(e[1]):8.1f} ms | DAC decode "
          f"{e[1].elapsed_time(e[2]):8.1f} ms | audio {tuple(audio.shape)} D2H {e[2].elapsed_time(e[3]):6.1f} ms")
