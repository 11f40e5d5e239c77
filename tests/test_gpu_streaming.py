"""Streaming audio out of generate(): the chunks an AudioChunkStreamer receives while the
CUDA-graph loop is still running concatenate to the audio generate() returns, which equals the one-shot result."""
import threading

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import parler_tts_b200 as ptts  # noqa: E402
from parler_tts_b200.streaming import chunk_schedule  # noqa: E402

from test_gpu_generate import DEV, _composite  # noqa: E402


def _inputs(B, seed=2):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(2, 100, (B, 7), generator=g).to(DEV), torch.randint(2, 100, (B, 4), generator=g).to(DEV)


def _long_model(max_length=56):
    model, dims, sd, dac_sd = _composite()
    model.generation_config.max_length = max_length  # TINY: max_position_embeddings 64, prompt 4
    return model


def test_streamed_chunks_equal_one_shot_audio():
    model = _long_model()
    ids, pids = _inputs(3)
    procs = [ptts.SuppressTokensAtOrAbove(32)]  # ids >= 32 are not codebook entries: every frame stays valid
    want = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs)
    got_chunks, pinned = [], []

    def on_chunk(first_frame, a):
        pinned.append(a.device.type == "cpu" and a.is_pinned())
        got_chunks.append((first_frame, a.clone()))

    streamer = ptts.AudioChunkStreamer(chunk_frames=16, on_chunk=on_chunk)
    audio = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs,
                           streamer=streamer)
    assert audio.shape == want.shape
    np.testing.assert_allclose(audio.cpu().numpy(), want.cpu().numpy(), atol=1e-5)
    F, hop = 56 - 4, 512
    sched = chunk_schedule(F, 16, 16)
    assert [f for f, _ in got_chunks] == [s[0] for s in sched]
    assert all(pinned)  # to_host=True: every chunk arrives in pinned host memory, copy completed
    cat = torch.cat([a for _, a in got_chunks], dim=1)
    assert cat.shape == (3, F * hop)
    np.testing.assert_allclose(cat.numpy(), want.cpu().numpy(), atol=1e-5)
    # the first chunk's de-delay + DAC decode went onto the second stream while decode steps were still being queued
    # (overlap, not a decode at the end); when its result is handed over depends on the GPU's timing
    assert streamer.stats[0]["decode_queued_at"] < 56 - 2
    assert streamer.stats[-1]["decode_queued_at"] is None  # the remainder is decoded after the loop
    # iterator protocol: the same chunks, then StopIteration
    assert [f for f, _ in streamer] == [s[0] for s in sched]


def test_streaming_from_a_consumer_thread_and_on_device():
    model = _long_model()
    ids, pids = _inputs(2, seed=5)
    procs = [ptts.SuppressTokensAtOrAbove(32)]
    want = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=True, top_k=8, seed=3, logits_processor=procs)
    streamer = ptts.AudioChunkStreamer(chunk_frames=16, to_host=False, timeout=60)
    got = []
    t = threading.Thread(target=lambda: got.extend((f, a.cpu()) for f, a in streamer))
    t.start()
    audio = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=True, top_k=8, seed=3,
                           logits_processor=procs, streamer=streamer)
    t.join(60)
    assert not t.is_alive()
    np.testing.assert_allclose(audio.cpu().numpy(), want.cpu().numpy(), atol=1e-5)
    np.testing.assert_allclose(torch.cat([a for _, a in got], dim=1).numpy(), want.cpu().numpy(), atol=1e-5)


def test_streaming_with_rows_that_finish_early():
    """No suppression: rows emit EOS / unused ids, the ragged path of modeling_parler_tts.py:2668-2677 applies.  What was
    streamed before the first invalid window is a prefix of every sample's final audio; the rest arrives in one piece."""
    model = _long_model()
    ids, pids = _inputs(3, seed=9)
    for seed in (1, 2, 3):
        kw = dict(input_ids=ids, prompt_input_ids=pids, do_sample=True, top_k=40, seed=seed)
        want = model.generate(**kw)
        chunks = []
        streamer = ptts.AudioChunkStreamer(chunk_frames=16, on_chunk=lambda f, a: chunks.append((f, a.clone())))
        audio = model.generate(streamer=streamer, **kw)
        assert audio.shape == want.shape
        np.testing.assert_allclose(audio.cpu().numpy(), want.cpu().numpy(), atol=1e-5)
        cat = torch.cat([a for _, a in chunks], dim=1)
        assert cat.shape == want.shape
        np.testing.assert_allclose(cat.numpy(), want.cpu().numpy(), atol=1e-5)


def test_streamer_falls_back_to_one_chunk_with_host_processors():
    """Python logits processors run the step-at-a-time host loop: no overlap, the audio arrives as one chunk."""
    model = _long_model(24)
    ids, pids = _inputs(2, seed=4)

    class Cut:
        def __call__(self, input_ids, scores):
            scores = scores.clone()
            scores[:, 32:] = -float("inf")
            return scores

    want = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=[Cut()])
    streamer = ptts.AudioChunkStreamer(chunk_frames=16)
    audio = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=[Cut()],
                           streamer=streamer)
    got = list(streamer)
    assert len(got) == 1 and got[0][0] == 0
    np.testing.assert_allclose(got[0][1].numpy(), want.cpu().numpy(), atol=1e-6)
    np.testing.assert_allclose(audio.cpu().numpy(), want.cpu().numpy(), atol=1e-6)


def test_decode_beside_the_loop_without_a_streamer_and_with_a_cta_limit():
    """generate() with model.overlap_decode_chunk_frames decodes finished chunks on the second stream while the loop runs
    (streaming.SilentChunks) and returns the same waveform; the CTA limit of the persistent conv kernels
    (ptts_dac_set_cta_limit, AudioChunkStreamer(decode_ctas=...)) changes the launch width only, not a sample."""
    from parler_tts_b200 import cabi

    model = _long_model()
    ids, pids = _inputs(3, seed=4)
    procs = [ptts.SuppressTokensAtOrAbove(32)]
    want = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs)
    model.overlap_decode_chunk_frames, model.overlap_decode_ctas = 16, 5
    got = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs)
    model.overlap_decode_chunk_frames, model.overlap_decode_ctas = 0, None
    assert got.shape == want.shape
    np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), atol=1e-5)
    lib = cabi.load()
    assert lib.ptts_dac_set_cta_limit(0) == 0  # the pipeline restored the default after its launches
    # the setter returns the previous value; a limited one-shot decode equals the unlimited one bit for bit
    codes = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs,
                           return_codes=True)
    eng = model.audio_encoder.engine(model.decoder.impl)
    full = eng.decode(codes)
    assert lib.ptts_dac_set_cta_limit(3) == 0
    try:
        narrow = eng.decode(codes)
    finally:
        assert lib.ptts_dac_set_cta_limit(0) == 3
    assert torch.equal(full, narrow)
    # a streamer with a CTA limit and the timeline trace switched on
    streamer = ptts.AudioChunkStreamer(chunk_frames=16, decode_ctas=7)
    streamer.trace = True
    audio = model.generate(input_ids=ids, prompt_input_ids=pids, do_sample=False, logits_processor=procs,
                           streamer=streamer)
    np.testing.assert_allclose(audio.cpu().numpy(), want.cpu().numpy(), atol=1e-5)
    tl = streamer.timeline
    assert len(tl["chunks"]) >= 1 and all(b >= a for _, _, _, a, b in tl["chunks"]) and tl["end_ms"] >= tl["loop_end_ms"]
