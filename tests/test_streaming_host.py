"""Host-side arithmetic of the streaming path (no GPU): which token columns a chunk needs, and that the chunk
schedule tiles the frames.  The de-delay geometry is checked against the oracle's restatement of
/root/reference/parler_tts/modeling_parler_tts.py:2633-2648."""
import numpy as np

from oracle import delay_pattern
from parler_tts_b200.streaming import chunk_ready_columns, chunk_schedule


def test_chunk_schedule_tiles_the_frames():
    for total, chunk, halo in ((860, 128, 16), (116, 16, 16), (40, 16, 16), (16, 16, 16), (1, 16, 16), (300, 64, 10)):
        sched = chunk_schedule(total, chunk, halo)
        assert sched[0][0] == 0 and sched[-1][1] == total
        for (f0, f1, lo, hi), nxt in zip(sched, sched[1:] + [None]):
            assert lo == max(0, f0 - halo) and lo <= f0 < f1 <= hi <= total
            if nxt is not None:
                assert f1 == nxt[0] and f1 - f0 == chunk and hi == f1 + halo  # full chunks see their whole right halo
            else:
                assert hi == total  # the last chunk ends at the true end (zero padding there, like the one-shot decode)


def test_window_is_complete_when_chunk_ready_columns_exist():
    """codes[b, k, f] = ids[b*K + k, f + k + 1]: with `chunk_ready_columns` columns generated, every frame of the window
    [lo, f1 + halo) only reads columns that already exist, and one frame more would not."""
    K, M, B = 9, 200, 2
    rng = np.random.default_rng(0)
    start = np.full((B * K, 1), 1025, dtype=np.int64)
    _, mask = delay_pattern.build_delay_pattern_mask(start, 1025, 1024, M, K)
    ids = rng.integers(0, 1024, size=(B * K, M)).astype(np.int64)
    applied = delay_pattern.apply_delay_pattern_mask(ids, mask)
    full = delay_pattern.revert_delay_pattern(ids, start, mask, 1025, 1024, K)  # (B, K, M - K)
    for f1, halo in ((16, 16), (64, 16), (128, 10)):
        cols = chunk_ready_columns(f1, halo, K)
        hi = f1 + halo
        need = max(f + k + 1 for f in range(hi) for k in range(K))
        assert need == cols - 1          # the last column read is cols - 1: exactly the generated ones
        assert max(hi + k + 1 for k in range(K)) == cols  # frame `hi` would need one more column
        win = np.stack([[applied[b * K + k, np.arange(hi) + k + 1] for k in range(K)] for b in range(B)])
        assert np.array_equal(win, full[:, :, :hi])


def test_dac_rates_are_read_from_the_checkpoint():
    """A DAC checkpoint states its own up-sampling rates (kernel of `decoder.model.{i}.block.1` = 2 x stride, descript
    DecoderBlock): DACModel(state_dict=...) must not assume the 44.1 kHz model's (8, 8, 4, 2)."""
    import pytest
    import torch

    import parler_tts_b200 as ptts
    from oracle import dac_oracle
    from parler_tts_b200.modeling import _dac_rates_of

    for rates in ((8, 8, 4, 2), (2, 4, 5, 8), (4, 4)):
        cfg = dict(n_codebooks=2, codebook_size=16, codebook_dim=4, latent_dim=32, decoder_dim=32, rates=rates)
        sd = dac_oracle.random_dac_state(cfg, seed=1)
        assert _dac_rates_of(sd) == rates
        m = ptts.DACModel(ptts.DACConfig(num_codebooks=2, codebook_size=16, latent_dim=32), state_dict=sd)
        assert m.rates == rates
        # an explicit argument still wins, random init keeps the reference model's rates
        assert ptts.DACModel(ptts.DACConfig(num_codebooks=2, codebook_size=16, latent_dim=32), state_dict=sd,
                             rates=rates[::-1]).rates == rates[::-1]
    assert ptts.DACModel(ptts.DACConfig(num_codebooks=2, codebook_size=16, latent_dim=32), decoder_dim=32,
                         codebook_dim=4).rates == (8, 8, 4, 2)
    with pytest.raises(ValueError):
        _dac_rates_of({"decoder.model.0.weight_v": torch.zeros(4, 4, 7)})


def test_generation_length_and_mode_rules_follow_the_reference():
    """Steps 1 and 6 of the reference generate (modeling_parler_tts.py:2416-2422, 2499-2521, 2575-2579): `max_new_tokens`
    wins over `max_length` and counts from the decoder input length, an unset `max_length` is the library default 20,
    `min_length > max_length` and beam search raise the reference's ValueErrors."""
    import pytest
    from transformers.generation.configuration_utils import GenerationConfig

    from parler_tts_b200.modeling import _final_max_length, _gcv, _resolve_generation

    base = GenerationConfig(max_length=2580, do_sample=True, pad_token_id=1024, eos_token_id=1024, bos_token_id=1025)
    gc, rest = _resolve_generation(base, None, dict(max_new_tokens=30, prompt_input_ids="kept for the model"), None)
    assert rest == {"prompt_input_ids": "kept for the model"} and base.max_new_tokens is None  # the caller's config is copied
    assert _final_max_length(gc, {}, input_len=1) == 31
    assert _final_max_length(_resolve_generation(base, None, dict(max_new_tokens=30), None)[0], {}, input_len=4) == 34
    assert _final_max_length(_resolve_generation(base, None, {}, None)[0], {}, 1) == 2580
    assert _final_max_length(_resolve_generation(base, None, dict(max_length=50), None)[0], {}, 1) == 50
    bare = GenerationConfig()
    bare.max_length = None
    assert _final_max_length(_resolve_generation(bare, None, {}, None)[0], {}, 1) == 20
    # an explicit generation_config replaces the model's; pad falls back to (the first) eos
    other = GenerationConfig(max_length=77, eos_token_id=[5, 6])
    gc, _ = _resolve_generation(base, other, {}, None)
    assert gc.max_length == 77 and gc.pad_token_id == 5
    # transformers <= 4.40 defaults for unset fields (top_k = 50 is the one that changes results)
    assert _gcv(gc, "top_k") == 50 and _gcv(gc, "temperature") == 1.0 and _gcv(gc, "num_beams") == 1
    with pytest.raises(ValueError, match="Unfeasible length constraints"):
        _final_max_length(_resolve_generation(base, None, dict(max_length=10, min_length=11), None)[0], {}, 1)
    for bad in (dict(num_beams=2), dict(num_beams=4, num_beam_groups=2, diversity_penalty=0.5)):
        with pytest.raises(ValueError, match="incompatible mode for generation"):
            _resolve_generation(base, None, bad, None)


def test_generation_plan_routes_every_field_or_raises():
    """`_Plan` = what the reference's `_get_logits_processor` / `_get_logits_warper` (modeling_parler_tts.py:2561-2567,
    2597) would build, split into the fused epilogue's native knobs and transformers' own classes run between the forward
    pass and the epilogue.  No field is dropped silently: it is native, or a python processor in the reference's order,
    or an error."""
    import pytest
    from transformers.generation import logits_process as lp
    from transformers.generation.configuration_utils import GenerationConfig

    from parler_tts_b200.modeling import _Plan

    def plan(user=(), cfg=False, T0=1, **kw):
        gc = GenerationConfig(pad_token_id=1024, eos_token_id=1024, bos_token_id=1025, max_length=50)
        gc.update(**kw)
        return _Plan(gc, T0, list(user), cfg, 1024)

    p = plan()  # greedy, nothing set
    assert not p.slow and not p.do_sample and p.rep_penalty == 1.0 and p.min_length == 0
    p = plan(min_new_tokens=5, T0=3, repetition_penalty=1.3, min_length=4)
    assert not p.slow and p.min_length == 8 and p.rep_penalty == 1.3  # native: max(min_length, T0 + min_new_tokens)
    p = plan(do_sample=True)
    assert not p.slow and (p.temperature, p.top_k, p.top_p) == (1.0, 50, 1.0)  # transformers <= 4.40 default top_k = 50
    p = plan(do_sample=True, temperature=0.7, top_k=20, top_p=0.9)
    assert not p.slow and (p.temperature, p.top_k, p.top_p) == (0.7, 20, 0.9)
    p = plan(do_sample=False, temperature=0.7, top_k=20, top_p=0.9)
    assert not p.slow and (p.temperature, p.top_k, p.top_p) == (1.0, 0, 1.0)  # warpers exist for sampling only
    # a warper the epilogue does not have: the whole warper chain runs as transformers' classes, in the library's order
    p = plan(do_sample=True, temperature=0.7, top_k=20, top_p=0.9, typical_p=0.5)
    assert p.slow and [type(w) for w in p.py_warpers] == [lp.TemperatureLogitsWarper, lp.TopKLogitsWarper,
                                                          lp.TopPLogitsWarper, lp.TypicalLogitsWarper]
    assert (p.temperature, p.top_k, p.top_p) == (1.0, 0, 1.0)  # and is not applied a second time natively
    # a processor the epilogue does not have: every default processor moves to python, reference order, user's list last
    user = [lambda ids, scores: scores]
    # (min_length would join them as MinLengthLogitsProcessor on the CUDA device: covered by tests/test_gpu_api.py)
    p = plan(user=user, no_repeat_ngram_size=2, repetition_penalty=1.2)
    assert p.slow and [type(x) for x in p.py_procs[:2]] == [lp.RepetitionPenaltyLogitsProcessor,
                                                           lp.NoRepeatNGramLogitsProcessor]
    assert p.py_procs[2] is user[0] and p.rep_penalty == 1.0 and p.min_length == 0
    assert plan(user=user, cfg=True).cfg_in_python and not plan(cfg=True).cfg_in_python
    assert plan(renormalize_logits=True).slow
    with pytest.raises(NotImplementedError):
        plan(encoder_repetition_penalty=1.5)
    with pytest.raises(ValueError):
        plan(do_sample=True, temperature=0.0)
    with pytest.raises(ValueError):
        plan(do_sample=True, top_p=0.0)
