"""Pins the oracle's restated token loop (oracle/generate_oracle.py: greedy selection, EOS freeze with PAD fill, early
stop, max_length stop, min_length / min_new_tokens, repetition penalty, the processor order) against the loop of the
INSTALLED `transformers` library itself - `GenerationMixin.generate` -> `_sample` - which is the third-party code the
reference's `generate` hands the decoder to (/root/reference/parler_tts/modeling_parler_tts.py:2582-2620).

The reference's composite `generate` cannot run under this transformers version, so the decoder is
replaced by a probe: a stateless logits function of (row, position, last applied token) with EOS boosts planted at
chosen (row, position) pairs, wrapped once as a `PreTrainedModel` whose `prepare_inputs_for_generation` applies the
delay-pattern mask to the history exactly as modeling_parler_tts.py:2107 does, and once as the oracle's `forward_fn`.
Both loops then see identical logits for identical histories; the raw sequences must be identical (int64, bit-exact),
including the columns after a row's EOS and the length at which the loop stops.

CPU only, live (no fixture): transformers is importable here and on the GPU box.
"""
import numpy as np
import pytest
import torch

from oracle import delay_pattern as dp
from oracle import generate_oracle
from oracle.decoder_oracle import DecoderDims

transformers = pytest.importorskip("transformers")
from transformers import GenerationMixin, PretrainedConfig, PreTrainedModel  # noqa: E402
from transformers.modeling_outputs import CausalLMOutput  # noqa: E402

K, V, PAD, EOS, BOS = 4, 40, 32, 32, 33
DIMS = DecoderDims(vocab_size=V, hidden_size=16, num_hidden_layers=1, num_attention_heads=1, ffn_dim=16, num_codebooks=K,
                   max_position_embeddings=64, pad_token_id=PAD, eos_token_id=EOS, bos_token_id=BOS)


class _ProbeConfig(PretrainedConfig):
    model_type = "ptts_loop_probe"


class _Probe(PreTrainedModel, GenerationMixin):
    """logits[r, t] = tanh(E[token] + Pos[t] + Row[r]) W + boost[r, t] * onehot(EOS): no state, no cache."""
    config_class = _ProbeConfig
    main_input_name = "input_ids"
    _supports_cache_class = False

    def __init__(self, config, rows, max_len, seed, eos_at, guided=False):
        super().__init__(config)
        self.guided = guided
        rows = 2 * rows if guided else rows  # the second half plays the unconditional rows
        g = torch.Generator().manual_seed(seed)
        D = 24
        self.E = torch.nn.Parameter(torch.randn(V + 2, D, generator=g), requires_grad=False)
        self.Pos = torch.nn.Parameter(torch.randn(max_len + 1, D, generator=g), requires_grad=False)
        self.Row = torch.nn.Parameter(torch.randn(rows, D, generator=g), requires_grad=False)
        self.W = torch.nn.Parameter(torch.randn(D, V, generator=g) * 2.0, requires_grad=False)
        boost = torch.zeros(rows, max_len + 1)
        for r, t in eos_at:
            boost[r, t:] = 100.0  # from position t on, row r wants EOS
        self.boost = torch.nn.Parameter(boost, requires_grad=False)
        self.pattern = None  # delay-pattern mask (rows, max_len), set by the caller
        self.seen = []       # histories the library handed to forward (after the mask), for the assertions below

    def logits_of(self, ids):
        ids = torch.as_tensor(ids, dtype=torch.int64)
        T = ids.shape[1]
        h = torch.tanh(self.E[ids] + self.Pos[:T][None] + self.Row[:, None])
        lo = h @ self.W
        lo[:, :, EOS] += self.boost[:, :T]
        lo[:, :, EOS + 1:] = -1e4  # ids above the EOS / PAD id (BOS, unused) never win
        return lo

    def forward(self, input_ids=None, attention_mask=None, **kwargs):
        self.seen.append(input_ids.clone())
        return CausalLMOutput(logits=self.logits_of(input_ids))

    def prepare_inputs_for_generation(self, input_ids, **kwargs):
        # modeling_parler_tts.py:2107: the mask is applied to the model's input, the loop keeps the raw tokens
        applied = torch.from_numpy(dp.apply_delay_pattern_mask(input_ids.numpy(), self.pattern))
        if self.guided:  # modeling_parler_tts.py:2108-2111: decoder args replicated across the batch dim
            applied = applied.repeat((2, 1))
        return {"input_ids": applied}


def _run_both(B, max_length, eos_at, seed, guidance_scale=None, **gen_kw):
    rows = B * K
    probe = _Probe(_ProbeConfig(pad_token_id=PAD, eos_token_id=EOS, bos_token_id=BOS), rows, max_length, seed, eos_at,
                   guided=guidance_scale is not None)
    probe.eval()
    start = np.full((rows, 1), BOS, dtype=np.int64)
    ids0, pattern = dp.build_delay_pattern_mask(start, BOS, PAD, max_length, K)
    probe.pattern = pattern
    assert ids0.shape == (rows, 1)

    # the library's loop (guidance as the reference wires it: its own batched processor appended to the caller's list,
    # generation_config.guidance_scale left unset, modeling_parler_tts.py:2555-2558)
    if guidance_scale is not None:
        from transformers import ClassifierFreeGuidanceLogitsProcessor, LogitsProcessorList
        gen_kw = dict(gen_kw, logits_processor=LogitsProcessorList([ClassifierFreeGuidanceLogitsProcessor(guidance_scale)]))
    with torch.no_grad():
        got = probe.generate(input_ids=torch.from_numpy(ids0), max_length=max_length, do_sample=False, num_beams=1,
                             use_cache=False, pad_token_id=PAD, eos_token_id=EOS, **gen_kw)
    got = got.numpy()

    # the oracle's loop over the same logits
    def fwd(inp, prompt, past):
        return probe.logits_of(inp), None

    enc = torch.zeros(B, 1, 1)
    okw = {k: v for k, v in gen_kw.items() if k in ("min_length", "min_new_tokens", "repetition_penalty")}
    want = generate_oracle.generate_codes(None, DIMS, enc, max_length=max_length, do_sample=False, forward_fn=fwd,
                                          guidance_scale=guidance_scale, **okw)
    return got, want, probe


def _assert_same(got, want):
    seq = want["sequences"]
    assert got.dtype == np.int64 and got.shape == seq.shape, (got.shape, seq.shape)
    assert np.array_equal(got, seq)


def test_runs_to_max_length_without_eos():
    got, want, probe = _run_both(B=2, max_length=14, eos_at=[], seed=1)
    _assert_same(got, want)
    assert got.shape[1] == 14 and want["steps"] == 13
    # the library fed the masked history (BOS / PAD triangles in place), not the raw tokens
    last = probe.seen[-1].numpy()
    assert np.array_equal(last, dp.apply_delay_pattern_mask(got[:, :last.shape[1]], probe.pattern))


def test_eos_freezes_a_row_and_pads_it_afterwards():
    # rows 1 and 6 finish early; the others run to max_length: frozen rows carry PAD from the column after their EOS
    got, want, _ = _run_both(B=2, max_length=16, eos_at=[(1, 4), (6, 7)], seed=2)
    _assert_same(got, want)
    assert got.shape[1] == 16
    for r, t in ((1, 4), (6, 7)):
        # the boost sits on input position t, whose logits choose column t + 1
        assert got[r, t + 1] == EOS
        col = int(np.argmax(got[r] == EOS))  # first EOS of the row (the planted one, or an earlier spontaneous one)
        assert 1 <= col <= t + 1 and (got[r, col:] == PAD).all()  # EOS id == PAD id in the Parler configs: all 32
    assert (got[0] != EOS).all()  # and a row that never chose EOS kept generating to max_length


def test_loop_stops_when_every_row_has_finished():
    eos_at = [(r, 3 + (r % 5)) for r in range(2 * K)]
    got, want, _ = _run_both(B=2, max_length=30, eos_at=eos_at, seed=3)
    _assert_same(got, want)
    assert got.shape[1] < 30  # early stop: unfinished.max() == 0
    assert got.shape[1] == 1 + want["steps"]


def test_min_length_and_min_new_tokens_hold_eos_back():
    eos_at = [(r, 2) for r in range(K)]
    got, want, _ = _run_both(B=1, max_length=20, eos_at=eos_at, seed=4, min_length=9)
    _assert_same(got, want)
    assert not (got[:, 1:9] == EOS).any() and (got[:, 9] == EOS).all()
    got, want, _ = _run_both(B=1, max_length=20, eos_at=eos_at, seed=4, min_new_tokens=6)
    _assert_same(got, want)
    assert not (got[:, 1:7] == EOS).any() and (got[:, 7] == EOS).all()


def test_repetition_penalty_sees_the_raw_history():
    got, want, _ = _run_both(B=2, max_length=18, eos_at=[(3, 9)], seed=5, repetition_penalty=1.7)
    _assert_same(got, want)
    plain, _, _ = _run_both(B=2, max_length=18, eos_at=[(3, 9)], seed=5)
    assert not np.array_equal(plain, got)  # the penalty changed decisions, so the comparison above exercised it


def test_final_codes_follow_from_the_library_sequences():
    """modeling_parler_tts.py:2633-2648: mask applied to the output ids, BOS / PAD columns dropped per codebook."""
    got, want, probe = _run_both(B=2, max_length=16, eos_at=[], seed=6)
    start = np.full((2 * K, 1), BOS, dtype=np.int64)
    codes = dp.revert_delay_pattern(got, start, probe.pattern, BOS, PAD, K)
    assert np.array_equal(codes, want["codes"])


def test_classifier_free_guidance_mixes_the_two_halves_like_the_library_processor():
    """modeling_parler_tts.py:2108-2111 + 2555-2558: 2R logit rows in, transformers' ClassifierFreeGuidanceLogitsProcessor
    mixes them to R score rows; EOS freeze and the stop rule act on the R rows."""
    got, want, probe = _run_both(B=2, max_length=16, eos_at=[(2, 5)], seed=7, guidance_scale=3.0)
    _assert_same(got, want)
    assert got.shape[0] == 2 * K and probe.seen[-1].shape[0] == 4 * K
    assert got[2, 6] == EOS and (got[2, 6:] == PAD).all()
    other, _, _ = _run_both(B=2, max_length=16, eos_at=[(2, 5)], seed=7, guidance_scale=1.5)
    assert not np.array_equal(other, got)  # the scale matters, so the mix was exercised


def test_sampling_candidate_set_is_the_library_warpers_set():
    """_get_logits_warper (modeling_parler_tts.py:2597): temperature, then top-k (ties with the k-th largest kept), then
    top-p.  The library's processed scores of the first step (`output_scores`) are finite exactly on the oracle's
    candidate set, and hold logits / temperature there; the draw itself (torch.multinomial) is not comparable."""
    from transformers.generation import logits_process as lp

    # class level, with ties at the k-th largest value
    g = torch.Generator().manual_seed(0)
    scores = torch.randn(5, V, generator=g)
    scores[:, 7] = scores[:, 3]
    scores[0, 3] = scores[0, 7] = scores[0].topk(4)[0][-1]  # row 0: a tie exactly at the k-th largest (k = 4)
    for k in (1, 4, 50):
        want = lp.TopKLogitsWarper(k)(None, scores.clone())
        for r in range(5):
            keep = generate_oracle.topk_candidates(scores[r].numpy(), k)
            assert np.array_equal(keep, torch.isfinite(want[r]).numpy()), (k, r)

    # loop level: what the library would sample from at step 0
    B, max_length, temp, k, p = 2, 2, 0.7, 6, 0.8
    rows = B * K
    probe = _Probe(_ProbeConfig(pad_token_id=PAD, eos_token_id=EOS, bos_token_id=BOS), rows, max_length, 11, [])
    probe.eval()
    start = np.full((rows, 1), BOS, dtype=np.int64)
    ids0, probe.pattern = dp.build_delay_pattern_mask(start, BOS, PAD, max_length, K)
    torch.manual_seed(0)
    with torch.no_grad():
        out = probe.generate(input_ids=torch.from_numpy(ids0), max_length=max_length, do_sample=True, temperature=temp,
                             top_k=k, top_p=p, use_cache=False, pad_token_id=PAD, eos_token_id=EOS, output_scores=True,
                             return_dict_in_generate=True)
    lib = out.scores[0]
    raw = probe.logits_of(torch.from_numpy(ids0))[:, -1].float()
    for r in range(rows):
        s = (raw[r].numpy() / np.float32(temp)).astype(np.float32)
        keep = generate_oracle.topp_candidates(s, generate_oracle.topk_candidates(s, k), p)
        fin = torch.isfinite(lib[r]).numpy()
        assert np.array_equal(keep, fin), r
        assert 1 <= keep.sum() <= k
        assert np.allclose(lib[r].numpy()[fin], s[fin], rtol=0, atol=1e-6)
        # and the oracle's draw can only land inside that set
        for u in (1e-6, 0.3, 0.999999):
            assert keep[generate_oracle.sample_row(raw[r].numpy(), u, temperature=temp, top_k=k, top_p=p)]
