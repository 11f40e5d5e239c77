#!/usr/bin/env python
"""Benchmark of the codec-token generation path (BASELINE.json: codec frames/s/GPU, Mini v0.1 architecture, bs=64,
greedy; ms / decode step).

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (one rank per GPU under torchrun for N>1)
    python bench.py --impl reference --steps K --warmup W    # the reference algorithm's CPU path (oracle port)

A "step" is ONE complete generation of the batch: prefill + (max_length - 2) cached decode steps (CUDA-graph
replays) + de-delay, i.e. B * N codec frames per rank.  `value` times it with the projected encoder states and prompt
ids already resident in HBM; `e2e` times the public `ParlerTTSForConditionalGeneration.generate()` call from host
token ids to host audio (H2D ids, T5 encode, projection, prefill, loop, de-delay, DAC decode, D2H audio).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MINI = dict(vocab_size=1088, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, ffn_dim=4096,
            num_codebooks=9, max_position_embeddings=4096, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
WORK = dict(B=64, S=48, P=32, frames=860)  # configs[1] shape, greedy (the metric's own wording)
T5 = dict(d_model=768, d_ff=2048, num_layers=12, num_heads=12)  # flan-t5-base shape
# BASELINE.json configs[3]: 36 layers x 1536 hidden (heads / ffn by the Mini rule d=64, ffn=4H), flan-t5-large-shaped
# encoder (--arch scaled; not the headline workload)
SCALED = dict(hidden_size=1536, num_hidden_layers=36, num_attention_heads=24, ffn_dim=6144)
T5_LARGE = dict(d_model=1024, d_ff=2816, num_layers=24, num_heads=16)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ---------------------------------------------------------------------------------------------- reference / CPU arm
def cpu_oracle_sample(n_decode_steps, threads=None, seed=0):
    """Times the oracle port of the reference path on the host cores: B=64 prefill + n cached decode steps of the
    Mini v0.1 architecture in fp32 (what `generate` does per step: apply-mask, forward, argmax, bookkeeping).
    Returns (frames_per_s over the decode steps, seconds, cores)."""
    import numpy as np
    import torch

    from oracle import decoder_oracle, delay_pattern

    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    dims = decoder_oracle.DecoderDims(**MINI)
    if not hasattr(cpu_oracle_sample, "_sd"):
        cpu_oracle_sample._sd = decoder_oracle.random_decoder_state(dims, seed=seed)
    sd = cpu_oracle_sample._sd
    g = torch.Generator().manual_seed(1234)
    B, S, P, K = WORK["B"], WORK["S"], WORK["P"], dims.num_codebooks
    enc = torch.randn(B, S, dims.hidden_size, generator=g) * 0.3
    prm = torch.randn(B, P, dims.hidden_size, generator=g) * 0.02
    start = np.full((B * K, 1), dims.bos_token_id, dtype=np.int64)
    ids, mask = delay_pattern.build_delay_pattern_mask(start, dims.bos_token_id, dims.pad_token_id,
                                                       WORK["frames"] + K, K)
    with torch.no_grad():
        logits, past = decoder_oracle.decoder_forward(sd, dims, torch.from_numpy(ids), enc, None, prm, None, None)
        t0 = time.perf_counter()
        for _ in range(n_decode_steps):
            tok = logits[:, -1, :1024].argmax(-1).numpy()
            ids = np.concatenate([ids, tok[:, None]], axis=1)
            inp = delay_pattern.apply_delay_pattern_mask(ids, mask)[:, -1:]
            logits, past = decoder_oracle.decoder_forward(sd, dims, torch.from_numpy(inp), enc, None, None, None, past)
        dt = time.perf_counter() - t0
    return B * n_decode_steps / dt, dt, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = 32
    for _ in range(min(args.warmup, 2)):
        cpu_oracle_sample(1)
    t0 = time.perf_counter()
    vals = [cpu_oracle_sample(n) for _ in range(args.steps)]
    wall = time.perf_counter() - t0
    fps = sum(WORK["B"] * n for _ in vals) / sum(v[1] for v in vals)
    cores = vals[0][2]
    sample = (f"oracle port (torch fp32 CPU restatement of the reference forward + delay-mask + argmax loop): "
              f"B={WORK['B']} prefill once, then {n} cached decode steps per timed step at context ~{WORK['P'] + 2}")
    line = {"impl": "reference", "metric": "codec frames/sec/GPU (Mini v0.1, bs=64, greedy)", "value": fps,
            "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * wall / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "Mini v0.1 architecture random-init, B=64, S=48, P=32, greedy; CPU sample",
                       "sample": sample},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------- CUDA arm
def build_models(device, dtype, seed=0):
    import torch
    from transformers import T5Config

    import parler_tts_b200 as ptts

    dec_cfg = ptts.ParlerTTSDecoderConfig(**MINI)
    t5 = T5Config(vocab_size=32128, d_kv=64, feed_forward_proj="gated-gelu", tie_word_embeddings=False, **T5)
    cfg = ptts.ParlerTTSConfig.from_sub_models_config(t5, ptts.DACConfig(), dec_cfg, vocab_size=32128)
    torch.manual_seed(seed)
    model = ptts.ParlerTTSForConditionalGeneration(cfg, seed=seed)
    model = model.to(device)
    model.decoder.to(dtype)
    model.text_encoder.to(dtype)
    return model


def run_cuda(args):
    import torch

    import parler_tts_b200 as ptts
    from parler_tts_b200 import cabi
    from parler_tts_b200.engine import DecoderEngine, GenState

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    lib = cabi.load()
    dtype = torch.float32 if args.dtype == "f32" else torch.bfloat16
    impl = cabi.IMPL_SIMPLE if args.kernels == "simple" else cabi.IMPL_FAST
    B, S, P, N = args.batch, WORK["S"], WORK["P"], args.frames
    K, M = MINI["num_codebooks"], args.frames + MINI["num_codebooks"]
    model = build_models(dev, dtype)
    model.decoder.impl = impl
    w = model.decoder.packed_weights()
    eng = DecoderEngine(w)
    g = torch.Generator().manual_seed(1234 + rank)
    input_ids_h = torch.randint(2, 32100, (B, S), generator=g).pin_memory()
    prompt_ids_h = torch.randint(2, 32100, (B, P), generator=g).pin_memory()
    procs = [ptts.SuppressTokensAtOrAbove(1024)]  # EOS-free deterministic run on random weights

    # ---- device-resident arm: projected encoder states + prompt ids already in HBM
    with torch.no_grad():
        ids_d = input_ids_h.to(dev)
        hs = model.encode_text(ids_d, torch.ones_like(ids_d))
        enc = model.project_encoder(hs, torch.ones_like(ids_d))
    prompt_ids = prompt_ids_h.to(dev)
    table = model.prompt_table()
    st = GenState(w, B, S, P, 1, M, impl=impl, suppress_from=1024, row_offset=rank * B * K,
                  do_sample=args.do_sample, top_k=50 if args.do_sample else 0)
    start = torch.full((B * K, 1), MINI["bos_token_id"], dtype=torch.int64, device=dev)
    codes = torch.empty(B, K, N, dtype=torch.int64, device=dev)
    from parler_tts_b200 import sharding
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    loop_ms = []

    def one_generation():
        st.reset(start, seed=0)
        eng.cross_kv(st, enc)
        eng.prefill(st, table, prompt_ids)
        e0, e1 = ev(), ev()
        e0.record()
        eng.run_loop(st, use_graph=not args.no_graph, poll_every=64)
        e1.record()
        cabi.call("ptts_dedelay", cabi.ptr(st.ids), M, cabi.ptr(st.mask), M, B, K, M, cabi.ptr(codes),
                  cabi.stream_ptr())
        if world > 1:  # the only collective of the path: gather the finished codes
            sharding.gather_padded(codes, MINI["pad_token_id"])
        return e0, e1

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 1)):
        one_generation()
    barrier()
    n0, r0 = lib.ptts_launch_count(), st.replays
    with ClockSampler(local) as clocks:
        t0, t1 = ev(), ev()
        t0.record()
        marks = [one_generation() for _ in range(args.steps)]
        t1.record()
        barrier()
    launches = int(lib.ptts_launch_count() - n0) + (st.replays - r0) * st.graph_launches
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    loop_ms = [a.elapsed_time(b) for a, b in marks]
    n_decode = M - 2  # forward passes after the prefill
    ms_decode = sum(loop_ms) / len(loop_ms) / n_decode
    value = world * B * N * args.steps / (total_ms / 1e3)
    assert int(st.read_ctrl()["stop_len"]) == M, st.read_ctrl()

    # ---- roofline of the dominant kernel (self-attention over the KV cache) timed live, same stream
    peak, peak_src = _peaks()
    heads, d, L, H = MINI["num_attention_heads"], 64, MINI["num_hidden_layers"], MINI["hidden_size"]
    es = 2 if dtype == torch.bfloat16 else 4
    ctx = P + M // 2
    ctx_t = torch.tensor([ctx - P], dtype=torch.int32, device=dev)
    reps = 5
    for timed in (False, True):
        a0, a1 = ev(), ev()
        a0.record()
        for _ in range(reps if timed else 1):
            for l in range(L):  # 24 distinct caches (24 x ~120 MB): nothing is served from L2 between launches
                cabi.call("ptts_attention", cabi.ptr(st.q), cabi.ptr(st.kv_self[l, 0]), cabi.ptr(st.kv_self[l, 1]),
                          cabi.dtype_enum(dtype), B, heads, d, 1, st.c_max, P, cabi.ptr(ctx_t), 0, None, 0,
                          cabi.ptr(st.attn), cabi.ptr(st.scratch), impl, cabi.stream_ptr())
        a1.record()
        torch.cuda.synchronize()
    attn_ms = a0.elapsed_time(a1) / (reps * L)
    attn_bytes = B * heads * (2 * ctx * d * es + 2 * d * es)
    achieved = attn_bytes / (attn_ms * 1e-3) / 1e9
    W = w.weight_elements()
    kvb = 2 * L * H * es
    avg_ctx = P + 1 + (M - 1) / 2.0
    step_bytes = W * es + B * kvb * (avg_ctx + S) + B * kvb + B * K * H * 4
    step_gbs = step_bytes / (ms_decode * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel at exactly these arguments, from the
    # committed `ncu --set full` capture profiles/r01_ncu_attn_full_v6.csv (122.31 MB read + 3.68 MB written)
    traffic = 127686656 if (B, N, es, heads) == (64, 860, 2, 16) and impl == cabi.IMPL_FAST else None
    roofline = {"bound": "hbm", "kernel": "self-attention decode (ptts_attention, ctx=%d)" % ctx, "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "profiles/r01_ncu_attn_full_v8.csv (ncu --set full, mean of 3 launches)",
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": attn_bytes, "ms_per_launch": attn_ms,
                "whole_step": {"algorithmic_bytes": step_bytes, "ms": ms_decode, "achieved": step_gbs,
                               "frac": step_gbs / peak}}

    # ---- end to end through the public API: host token ids -> generate() -> host audio
    model.generation_config.max_length = M
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    audio_h = None

    def one_e2e():
        nonlocal audio_h
        i_d = input_ids_h.to(dev, non_blocking=True)
        p_d = prompt_ids_h.to(dev, non_blocking=True)
        out = model.generate(input_ids=i_d, prompt_input_ids=p_d, do_sample=args.do_sample, logits_processor=procs, seed=0,
                             row_offset=rank * B)
        if audio_h is None:
            audio_h = torch.empty(out.shape, dtype=out.dtype).pin_memory()
        audio_h.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    e2e = None
    if not args.skip_e2e:
        del st
        torch.cuda.empty_cache()
        one_e2e()
        barrier()
        c0, c1 = ev(), ev()
        c0.record()
        for _ in range(e2e_steps):
            one_e2e()
        c1.record()
        barrier()
        ems = torch.tensor([c0.elapsed_time(c1)], device=dev)
        if world > 1:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
        e2e = {"value": world * B * N * e2e_steps / (float(ems.item()) / 1e3), "unit": "frames/s",
               "h2d_bytes_per_step": int(input_ids_h.numel() * 8 + prompt_ids_h.numel() * 8),
               "d2h_bytes_per_step": int(audio_h.numel() * 4), "steps": e2e_steps,
               "includes": "H2D ids, T5 encode, enc_to_dec_proj, prefill, decode loop, de-delay, DAC decode, D2H audio"}

    cpu = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        fps, secs, cores = cpu_oracle_sample(args.cpu_steps)
        cpu = {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
               "sample": f"oracle port, fp32, B={WORK['B']}: {args.cpu_steps} cached decode steps after one prefill "
                         f"({secs:.1f} s), context ~{P + 2}"}
    if rank == 0:
        line = {"metric": "codec frames/sec/GPU (Mini v0.1, bs=64, greedy)", "value": value, "unit": "frames/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
                "ms_per_decode_step": ms_decode, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16" if dtype == torch.bfloat16 else "f32", "data": "synthetic",
                "config": {"workload": f"{'Mini v0.1' if args.arch == 'mini' else 'scaled'} architecture random-init "
                                       f"({L}x{H}, 9 codebooks), B={B} per GPU, "
                                       f"S={S}, P={P}, {N} frames (max_length {M}), "
                                       f"{'top-k 50 sampling' if args.do_sample else 'greedy'}, ids>=1024 suppressed, "
                                       "static KV cache + CUDA graph",
                           "kernels": args.kernels, "kv_dtype": "bf16" if dtype == torch.bfloat16 else "f32",
                           "l2": f"per-step working set (weights {W * es / 1e6:.0f} MB + KV) exceeds the 126 MB L2; no flush needed",
                           "parallelism": f"batch-sharded replicas x{world}, NCCL all_gather of codes only"},
                "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--kernels", default="fast", choices=["fast", "simple"])
    ap.add_argument("--batch", type=int, default=WORK["B"])
    ap.add_argument("--frames", type=int, default=WORK["frames"])
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-steps", type=int, default=128)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--do-sample", action="store_true", help="top-k 50 sampling instead of greedy (configs[1] wording)")
    ap.add_argument("--arch", default="mini", choices=["mini", "scaled"])
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    args = ap.parse_args()
    if args.arch == "scaled":
        MINI.update(SCALED)
        T5.update(T5_LARGE)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
