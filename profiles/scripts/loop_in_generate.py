"""Where the decode loop's time inside generate() differs from the device-resident loop of bench.py (1.203 ms per step):
generate(return_codes=True) (T5 + cross K/V + prefill + loop, no codec) against the full call, alternating, with the SM
clock sampled during each call (Mini v0.1, B=64, 860 frames, bf16)."""
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import parler_tts_b200 as ptts  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, P, N = 64, 48, 32, 860
model = bench.build_models(dev, torch.bfloat16)
model.generation_config.max_length = N + 9
g = torch.Generator().manual_seed(0)
ids = torch.randint(2, 32100, (B, S), generator=g).to(dev)
pids = torch.randint(2, 32100, (B, P), generator=g).to(dev)
procs = [ptts.SuppressTokensAtOrAbove(1024)]
# "ragged": descriptions of 40 .. 48 tokens (a mask with zeros: the masked cross-attention instantiation runs)
RAGGED = "ragged" in sys.argv[1:]
am = torch.ones(B, S, dtype=torch.int64, device=dev)
if RAGGED:
    for b in range(B):
        am[b, 40 + b % 9:] = 0


class Clock:
    def __enter__(self):
        self.samples, self.stop = [], False
        self.t = threading.Thread(target=self.run)
        self.t.start()
        return self

    def run(self):
        while not self.stop:
            self.samples.append(torch.cuda.clock_rate(0))
            time.sleep(0.02)

    def __exit__(self, *a):
        self.stop = True
        self.t.join()


def run(codes_only):
    torch.cuda.synchronize()
    with Clock() as c:
        t0 = time.perf_counter()
        out = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs, seed=0,
                             return_codes=codes_only, attention_mask=am if RAGGED else None)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    s = sorted(c.samples)
    return ms, s[len(s) // 2] if s else 0, s[0] if s else 0


for _ in range(2):
    run(True)
    run(False)
print("ragged description mask" if RAGGED else "no mask (all ones)")
seq = (("codes only", True),) * 4 if RAGGED else (("codes only", True), ("codes only", True), ("full call ", False), ("codes only", True),
                                                  ("full call ", False), ("full call ", False), ("codes only", True), ("codes only", True))
for label, co in seq:
    ms, med, lo = run(co)
    print(f"{label}: {ms:8.1f} ms   SM clock median {med} MHz, min {lo} MHz")
