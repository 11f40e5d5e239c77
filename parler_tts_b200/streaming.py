"""Streaming audio output of generate(): incremental de-delay + chunked DAC decode that
overlaps the token loop.

The reference only offers the token-level hook `streamer.put(token_ids)` (/root/reference/parler_tts/
modeling_parler_tts.py:2540-2541 and the transformers loop); turning tokens into sound while the loop still runs -
undo the delay pattern for the frames that are complete, decode them with the codec, hand the samples on - is left to
the caller.  Here that consumer is part of the path: pass an `AudioChunkStreamer` as `streamer=` and generate() keeps its
CUDA-graph loop (no per-step host sync) and, every `chunk_frames` frames, enqueues on a second stream

    ptts_dedelay (window of the token matrix) -> validity flag -> DAC decode with a halo of real context -> D2H copy

while the loop's stream goes on replaying decode steps.  `put_audio(first_frame, samples)` is called as chunks
complete; the concatenation of the chunks equals the audio generate() returns (and generate() still returns it).

Frame f of the de-delayed codes is complete once token column f + K exists (codes[b, k, f] = ids[b*K + k, f + k + 1],
modeling_parler_tts.py:2633-2648), so the chunk [f0, f1) with `halo` frames of right context can be decoded as soon as
f1 + halo + K columns have been generated.
"""
import queue

import torch

from . import cabi


def chunk_ready_columns(f1, halo, num_codebooks, total_frames=None):
    """Token columns that must exist before frames [.., f1) plus `halo` frames of right context are complete."""
    hi = f1 + halo if total_frames is None else min(f1 + halo, total_frames)
    return hi + num_codebooks


def chunk_schedule(total_frames, chunk_frames, halo):
    """[(f0, f1, lo, hi)]: the chunks a run of `total_frames` frames is streamed in and the window [lo, hi) each one is
    decoded from.  Every chunk but the last is `chunk_frames` long and sees `halo` real frames on both sides (fewer only
    at the true start); the last one takes the remainder and ends at the true end."""
    out, f0 = [], 0
    while total_frames - f0 >= chunk_frames + halo:  # a full chunk only if its right halo exists too
        f1 = f0 + chunk_frames
        out.append((f0, f1, max(0, f0 - halo), f1 + halo))
        f0 = f1
    if f0 < total_frames or not out:
        out.append((f0, total_frames, max(0, f0 - halo), total_frames))
    return out


class AudioChunkStreamer:
    """Receives audio while generate() is still producing tokens.

    `put_audio(first_frame, audio)` is called from the generate() thread with `audio` (B, n_frames * hop) fp32 - a
    pinned host tensor (`to_host=True`, the copy has completed) or a device tensor whose producing work has completed.
    Chunks arrive in order; `end()` is called after the last one.  Iterating over the streamer from another thread
    yields the same `(first_frame, audio)` pairs (like transformers' TextIteratorStreamer).  `put(token_ids)` - the
    reference's token-level hook - is accepted and ignored, so the object is also a valid BaseStreamer."""

    def __init__(self, chunk_frames=128, on_chunk=None, to_host=True, timeout=None, decode_ctas=None):
        if chunk_frames < 16:
            raise ValueError("chunk_frames must be at least 16")
        self.chunk_frames = int(chunk_frames)
        # CTAs the chunk decodes that run beside the token loop may occupy (None = one per SM): the codec's conv kernels
        # are persistent, one CTA per SM, and with all SMs taken the loop's kernels queue behind every conv launch
        self.decode_ctas = None if decode_ctas is None else int(decode_ctas)
        self.on_chunk = on_chunk
        self.to_host = bool(to_host)
        self.timeout = timeout
        self.queue = queue.Queue()
        # one dict per chunk: first_frame, samples, steps_enqueued (decode steps the host had queued when the chunk was
        # handed over), decode_queued_at (decode steps queued when the chunk's de-delay + DAC decode were enqueued on
        # the second stream; None for the remainder decoded after the loop)
        self.stats = []
        self._stop = object()

    def put(self, value):  # token ids of the reference's streamer protocol: not needed, the path de-delays on the device
        pass

    def put_audio(self, first_frame, audio, steps_enqueued=None, decode_queued_at=None):
        self.stats.append(dict(first_frame=int(first_frame), samples=int(audio.shape[-1]),
                               steps_enqueued=steps_enqueued, decode_queued_at=decode_queued_at))
        if self.on_chunk is not None:
            self.on_chunk(first_frame, audio)
        self.queue.put((first_frame, audio))

    def end(self):
        self.queue.put(self._stop)

    def __iter__(self):
        return self

    def __next__(self):
        item = self.queue.get(timeout=self.timeout)
        if item is self._stop:
            raise StopIteration
        return item


class SilentChunks(AudioChunkStreamer):
    """generate()'s own consumer when no streamer was passed: the chunks stay on the device and only the assembled
    waveform is returned - the codec work then overlaps the token loop instead of following it."""

    def __init__(self, chunk_frames, decode_ctas=None):
        super().__init__(chunk_frames=chunk_frames, to_host=False, decode_ctas=decode_ctas)

    def put_audio(self, first_frame, audio, steps_enqueued=None, decode_queued_at=None):
        pass

    def end(self):
        pass


class ChunkPipeline:
    """Host side of the overlapped chunk decode for one generate() call (no audio prompt: T0 == 1)."""

    def __init__(self, dac_engine, st, streamer, num_codebooks, halo=None):
        self.dac, self.st, self.streamer, self.K = dac_engine, st, streamer, int(num_codebooks)
        self.halo = dac_engine.STREAM_HALO if halo is None else int(halo)
        self.chunk = streamer.chunk_frames
        hop = 1
        for r in dac_engine.rates:
            hop *= r
        self.hop = hop
        self.F_max = st.max_len - self.K  # frames of a run that goes to max_length
        self.main = torch.cuda.current_stream()
        self.side = torch.cuda.Stream()
        self.audio = torch.zeros(st.Bu, max(self.F_max, 1) * hop, device=st.ids.device, dtype=torch.float32)
        self.audio.record_stream(self.side)
        self.next_f = 0       # first frame not yet handed to the side stream
        self.emitted_f = 0    # first frame not yet handed to the streamer
        self.stopped = False  # a window held an id outside the codebook (EOS / PAD / BOS / unused): the rest is decoded at the end
        self.pending = []
        self.steps_enqueued = 0
        # timeline for profiles/scripts/overlap_time.py: events with timing only when the streamer asks (trace = True)
        self.trace = [] if getattr(streamer, "trace", False) else None
        if self.trace is not None:
            self.t_ref = torch.cuda.Event(enable_timing=True)
            self.t_ref.record(self.main)

    # ---- called by the loop after every batch of enqueued decode steps
    def progress(self, done_steps):
        self.steps_enqueued = done_steps
        len_enq = self.st.T0 + 1 + done_steps  # columns that exist once the queued steps have run (if nothing stopped)
        while not self.stopped:
            f0, f1 = self.next_f, self.next_f + self.chunk
            if f1 + self.halo > self.F_max or len_enq < chunk_ready_columns(f1, self.halo, self.K):
                break
            self._enqueue(f0, f1, max(0, f0 - self.halo), f1 + self.halo)
            self.next_f = f1
        self._poll(block=False)

    def _decode_window(self, lo, hi):
        """De-delayed codes of frames [lo, hi) (window of the live token matrix) and their waveform; current stream."""
        st, K = self.st, self.K
        win = torch.empty(st.Bu, K, hi - lo, device=st.ids.device, dtype=torch.int64)
        cabi.call("ptts_dedelay", cabi.ptr(st.ids[:, lo:]), st.max_len, cabi.ptr(st.mask[:, lo:]), st.max_len, st.Bu, K,
                  (hi - lo) + K, cabi.ptr(win), cabi.stream_ptr())
        return win

    def _enqueue(self, f0, f1, lo, hi):
        st, hop = self.st, self.hop
        ready = torch.cuda.Event()
        ready.record(self.main)
        with torch.cuda.stream(self.side):
            self.side.wait_event(ready)
            if self.trace is not None:
                t_a = torch.cuda.Event(enable_timing=True)
                t_a.record(self.side)
            win = self._decode_window(lo, hi)
            bad = ((win < 0) | (win >= self.dac.codebook_size)).any() | (st.ctrl[0] < hi + self.K)
            lib, limit = cabi.load(), self.streamer.decode_ctas
            prev = lib.ptts_dac_set_cta_limit(limit) if limit else 0
            try:
                wav = self.dac.decode(win)
            finally:
                if limit:
                    lib.ptts_dac_set_cta_limit(prev)
            piece = wav[:, 0, (f0 - lo) * hop : (f1 - lo) * hop]
            self.audio[:, f0 * hop : f1 * hop].copy_(piece)
            host = None
            if self.streamer.to_host:
                host = torch.empty(piece.shape, dtype=torch.float32, pin_memory=True)
                host.copy_(piece, non_blocking=True)
            flag = torch.empty(1, dtype=torch.int32, pin_memory=True)
            flag.copy_(bad.to(torch.int32).reshape(1), non_blocking=True)
            done = torch.cuda.Event(enable_timing=self.trace is not None)
            done.record(self.side)
            if self.trace is not None:
                self.trace.append((f0, f1, self.steps_enqueued, t_a, done))
        self.pending.append((f0, f1, host, flag, done, self.steps_enqueued))

    def _poll(self, block):
        while self.pending:
            f0, f1, host, flag, done, queued_at = self.pending[0]
            if block:
                done.synchronize()
            elif not done.query():
                return
            self.pending.pop(0)
            if self.stopped:
                continue
            if int(flag[0]) != 0:
                self.stopped = True  # later windows are dropped; finish() decodes the remainder the reference's way
                continue
            out = host if host is not None else self.audio[:, f0 * self.hop : f1 * self.hop]
            self.streamer.put_audio(f0, out, self.steps_enqueued, queued_at)
            self.emitted_f = f1

    # ---- after the loop
    def finish(self, codes, decode_generated):
        """codes (Bu, K, F): the final de-delayed codes.  Returns the full audio (Bu, T) like the one-shot path and hands
        the part not streamed yet to the streamer."""
        if self.trace is not None:
            self.t_loop_end = torch.cuda.Event(enable_timing=True)
            self.t_loop_end.record(self.main)
        self._poll(block=True)
        hop, F = self.hop, codes.shape[2]
        tail_from = self.emitted_f
        special = bool(((codes < 0) | (codes >= self.dac.codebook_size)).any().item()) if codes.numel() else False
        if not special and F > 0:
            # every frame is a valid code: the remainder [tail_from, F) is one more window with left context
            self.main.wait_stream(self.side)
            lo = max(0, tail_from - self.halo)
            wav = self.dac.decode(codes[:, :, lo:F].contiguous())
            if self.audio.shape[1] < F * hop:
                grown = torch.zeros(codes.shape[0], F * hop, device=codes.device, dtype=torch.float32)
                grown[:, : tail_from * hop] = self.audio[:, : tail_from * hop]
                self.audio = grown
            self.audio[:, tail_from * hop : F * hop] = wav[:, 0, (tail_from - lo) * hop :]
            audio = self.audio[:, : F * hop]
        else:
            # ids outside the codebook survived the de-delay: the reference drops those frames per sample and pads
            # (modeling_parler_tts.py:2668-2677).  What was streamed is a prefix of every sample's result (all frames of
            # a streamed window, halo included, were valid for every sample), the rest follows in one piece.
            self.main.wait_stream(self.side)
            audio = decode_generated(codes)
        rest = audio[:, tail_from * hop :]
        if rest.shape[1] > 0 or tail_from == 0:
            out = rest
            if self.streamer.to_host:
                out = torch.empty(rest.shape, dtype=torch.float32, pin_memory=True)
                out.copy_(rest, non_blocking=False)
            self.streamer.put_audio(tail_from, out, self.steps_enqueued)
        self.streamer.end()
        if self.trace is not None:
            t_end = torch.cuda.Event(enable_timing=True)
            t_end.record(self.main)
            t_end.synchronize()
            self.streamer.timeline = dict(
                loop_end_ms=self.t_ref.elapsed_time(self.t_loop_end), end_ms=self.t_ref.elapsed_time(t_end),
                chunks=[(f0, f1, steps, self.t_ref.elapsed_time(a), self.t_ref.elapsed_time(b))
                        for f0, f1, steps, a, b in self.trace])
        return audio
