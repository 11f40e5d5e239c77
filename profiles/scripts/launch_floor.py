"""Measures the per-kernel latency floor of a CUDA-graph replay on this box: chains of 267 dependent trivial kernels
(the decode step's launch count), empty or with one load+store, with and without programmatic dependent launch
(run twice: PTTS_PDL=1 / PTTS_PDL=0)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from parler_tts_b200 import cabi  # noqa: E402

torch.cuda.set_device(0)
buf = torch.zeros(148 * 128, device="cuda")
N = 267
for mode, ctas in ((0, 1), (0, 148), (1, 64), (1, 148)):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        cabi.call("ptts_diag_chain", N, mode, ctas, cabi.ptr(buf), cabi.stream_ptr())
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            cabi.call("ptts_diag_chain", N, mode, ctas, cabi.ptr(buf), cabi.stream_ptr())
        for _ in range(5):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 50 / N
        print(f"PDL={os.environ.get('PTTS_PDL', '1')} mode={mode} ctas={ctas}: {us:.2f} us per kernel in a {N}-node graph")
