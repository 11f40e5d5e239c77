"""ORACLE (test infrastructure, CPU) - the autoregressive codec-token loop and its logits epilogue.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this.
It is the checker for the CUDA path, never the product path.

Restates
  * ParlerTTSForConditionalGeneration.generate glue     /root/reference/parler_tts/modeling_parler_tts.py:2416-2648
  * prepare_inputs_for_generation                        modeling_parler_tts.py:2081-2148
  * _prepare_decoder_input_ids_for_generation            modeling_parler_tts.py:2150-2195
  * greedy_search / sample of the third-party dependency `transformers` (pinned only as >=4.34.0 in
    /root/reference/setup.py:20; docs link v4.39.3, /root/reference/training/README.md:19), whose published
    algorithm per row is: scores = processors(logits[:, -1]); [warpers: temperature, top-k keeps scores >= k-th
    largest]; token = argmax (first index on ties) | multinomial(softmax); token = token*unfinished +
    pad*(1-unfinished); append; unfinished &= token != eos; stop when unfinished.max()==0 or len >= max_length.
    Call sites: modeling_parler_tts.py:2582-2593 (greedy), 2597-2620 (sample).
  * ClassifierFreeGuidanceLogitsProcessor (transformers): log_softmax both halves, uncond + s*(cond-uncond).

`torch.multinomial`'s stream cannot be reproduced on another device; sampling parity is therefore defined against
this file's inverse-CDF draw over a Philox4x32-10 uniform keyed by (seed, row, step) - the same construction the CUDA
epilogue uses ("parity unpinned" for the random stream itself; the candidate set, temperature and renormalisation
follow the library's published algorithm).

The loop itself is pinned twice: against a probe that drives the reference decoder (tests/golden/loop_tiny.npz), and,
live, against the installed transformers library's own `GenerationMixin.generate` / `_sample` loop over a stateless
probe model (tests/test_oracle_vs_transformers_loop.py: EOS freeze + PAD fill, early stop, max_length stop, min_length /
min_new_tokens, repetition penalty - raw sequences bit-identical).
"""
import numpy as np
import torch

from . import delay_pattern as dp
from .decoder_oracle import DecoderDims, decoder_forward

# ---------------------------------------------------------------------------------------------- Philox4x32-10
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32(counter, key):
    """counter: (...,4) uint32, key: (...,2) uint32 -> (...,4) uint32 (10 rounds)."""
    c = np.array(counter, dtype=np.uint32, copy=True)
    k = np.array(key, dtype=np.uint32, copy=True)
    for _ in range(10):
        p0 = _M0 * c[..., 0].astype(np.uint64)
        p1 = _M1 * c[..., 2].astype(np.uint64)
        hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
        hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
        c = np.stack([hi1 ^ c[..., 1] ^ k[..., 0], lo1, hi0 ^ c[..., 3] ^ k[..., 1], lo0], axis=-1)
        with np.errstate(over="ignore"):
            k = np.stack([k[..., 0] + _W0, k[..., 1] + _W1], axis=-1)
    return c


def uniform_for(seed, row, step):
    """The uniform in (0,1) that decision (global row, step) consumes: Philox(counter=(row, step, 0, 0),
    key=(seed_lo, seed_hi)) word 0 -> (x + 0.5) * 2^-32 in fp32 arithmetic."""
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    ctr = np.array([row & 0xFFFFFFFF, step & 0xFFFFFFFF, 0, 0], dtype=np.uint32)
    key = np.array([seed & 0xFFFFFFFF, seed >> 32], dtype=np.uint32)
    w = philox4x32(ctr, key)[0]
    return np.float32((np.float64(w) + 0.5) * (2.0 ** -32))


# ---------------------------------------------------------------------------------------------- logits epilogue
def classifier_free_guidance(scores, guidance_scale):
    """scores (2R, V): first half conditional, second half unconditional."""
    r = scores.shape[0] // 2
    lp = torch.log_softmax(scores, dim=-1)
    cond, unc = lp[:r], lp[r:]
    return unc + guidance_scale * (cond - unc)


def topk_candidates(scores_row, top_k):
    """TopKLogitsWarper: keep every score >= the k-th largest (ties kept)."""
    k = min(max(int(top_k), 1), scores_row.shape[-1])
    kth = np.sort(scores_row)[-k]
    return scores_row >= kth


def topp_candidates(scores_row, keep, top_p):
    """TopPLogitsWarper (min_tokens_to_keep=1) over the candidates `keep` (the others are already -inf):
    sort ascending, cumulative softmax probabilities, remove those with cumsum <= 1 - top_p."""
    s = np.where(keep, scores_row, -np.inf).astype(np.float32)
    order = np.argsort(s, kind="stable")
    ss = s[order]
    m = ss[-1]
    p = np.exp((ss - m).astype(np.float32)).astype(np.float32)
    p = p / p.sum(dtype=np.float32)
    cum = np.cumsum(p, dtype=np.float32)
    remove_sorted = cum <= np.float32(1.0 - top_p)
    remove_sorted[-1:] = False
    remove = np.zeros_like(remove_sorted)
    remove[order] = remove_sorted
    return keep & ~remove


def repetition_penalty_(scores, ids, penalty):
    """RepetitionPenaltyLogitsProcessor: score[t] = score[t] * p if < 0 else / p for every t in the row's raw history.
    With guidance the scores hold 2R rows and ids R rows: only the first R (conditional) rows are touched, which is what
    the processor's gather / scatter over mismatched batch sizes does."""
    s = scores.clone()
    idx = torch.from_numpy(np.asarray(ids, dtype=np.int64))
    R = idx.shape[0]
    g = torch.gather(s[:R], 1, idx)
    g = torch.where(g < 0, g * penalty, g / penalty)
    s[:R] = s[:R].scatter(1, idx, g)
    return s


def sample_row(scores_row, u, temperature=1.0, top_k=50, top_p=1.0):
    """Inverse-CDF draw in index order over the kept set, fp32 arithmetic in a fixed order:
    p_i = exp(s_i - max) (kept, else 0); walk i ascending accumulating p_i; first i with cumsum > u * total."""
    s = np.asarray(scores_row, dtype=np.float32)
    if temperature != 1.0:
        s = (s / np.float32(temperature)).astype(np.float32)
    keep = topk_candidates(s, top_k) if top_k and top_k > 0 else np.isfinite(s) | (s == np.inf)
    if top_p is not None and top_p < 1.0:
        keep = topp_candidates(s, keep, top_p)
    m = s[keep].max()
    p = np.where(keep, np.exp((s - m).astype(np.float32)).astype(np.float32), np.float32(0.0)).astype(np.float32)
    idx = np.nonzero(keep)[0]
    total = np.float32(0.0)
    for i in idx:  # sequential fp32 sum in ascending index order (matches the CUDA epilogue)
        total = np.float32(total + p[i])
    thresh = np.float32(np.float32(u) * total)
    acc = np.float32(0.0)
    for i in idx:
        acc = np.float32(acc + p[i])
        if acc > thresh:
            return int(i)
    return int(idx[-1])


def select_tokens(scores, do_sample, seed=0, step=0, row_offset=0, temperature=1.0, top_k=50, top_p=1.0):
    """scores (R, V) torch -> (R,) int64 numpy."""
    s = scores.detach().to(torch.float32).cpu().numpy()
    if not do_sample:
        return s.argmax(axis=-1).astype(np.int64)  # numpy argmax = first maximal index, like torch
    out = np.empty(s.shape[0], dtype=np.int64)
    for r in range(s.shape[0]):
        u = uniform_for(seed, row_offset + r, step)
        out[r] = sample_row(s[r], u, temperature, top_k, top_p)
    return out


# ---------------------------------------------------------------------------------------------- the loop
def generate_codes(
    sd,
    dims: DecoderDims,
    encoder_hidden_states,
    encoder_attention_mask=None,
    prompt_hidden_states=None,
    prompt_attention_mask=None,
    max_length=20,
    do_sample=False,
    temperature=1.0,
    top_k=50,
    seed=0,
    row_offset=0,
    guidance_scale=None,
    decoder_input_ids=None,
    suppress_from=None,
    forward_fn=None,
    return_trace=False,
    top_p=1.0,
    repetition_penalty=1.0,
    min_length=0,
    min_new_tokens=None,
    logits_processors=None,
):
    """Runs the loop.  Returns dict(sequences (B*K, Lout) raw ids incl. the start column, pattern_mask, codes (B,K,F),
    steps, and optionally per-step logits).  `suppress_from`: if set, logits at vocabulary index >= suppress_from are
    -inf'ed before selection (a user logits_processor; modeling_parler_tts.py:2347,2561-2567).
    `forward_fn(input_ids, prompt_hidden_states, past)` can replace the oracle forward (used to drive the reference
    decoder when generating fixtures)."""
    K = dims.num_codebooks
    B = encoder_hidden_states.shape[0] if encoder_hidden_states is not None else prompt_hidden_states.shape[0]
    bos, pad, eos = dims.bos_token_id, dims.pad_token_id, dims.eos_token_id
    cfg = guidance_scale is not None and guidance_scale > 1

    start = np.full((B * K, 1), bos, dtype=np.int64)  # modeling_parler_tts.py:2174-2181
    if decoder_input_ids is not None:
        user = np.asarray(decoder_input_ids, dtype=np.int64)
        if (user[..., 0] != bos).all():  # modeling_parler_tts.py:2185-2186
            user = np.concatenate([start, user], axis=-1)
        start = user
    ids, pattern = dp.build_delay_pattern_mask(start, bos, pad, max_length, K)

    enc, enc_mask, prm, prm_mask = encoder_hidden_states, encoder_attention_mask, prompt_hidden_states, prompt_attention_mask
    if cfg:  # modeling_parler_tts.py:2231-2237 (null text conditioning) and 2115-2118 (prompt repeated)
        enc = torch.cat([enc, torch.zeros_like(enc)], dim=0)
        if enc_mask is not None:
            enc_mask = torch.cat([enc_mask, torch.zeros_like(enc_mask)], dim=0)
        if prm is not None:
            prm = prm.repeat(2, 1, 1)
        if prm_mask is not None:
            prm_mask = prm_mask.repeat(2, 1)

    if forward_fn is None:
        def forward_fn(inp, prompt, past):
            return decoder_forward(sd, dims, inp, enc, enc_mask, prompt, prm_mask, past)

    # MinLength / MinNewTokensLength processors (prompt_length_to_skip = the start columns) as one column threshold
    min_cols = int(min_length or 0)
    if min_new_tokens:
        min_cols = max(min_cols, start.shape[1] + int(min_new_tokens))
    unfinished = np.ones(B * K, dtype=np.int64)
    past = None
    trace = []
    step = 0
    while True:
        applied = dp.apply_delay_pattern_mask(ids, pattern)  # whole history, modeling_parler_tts.py:2107
        model_in = applied if past is None else applied[:, -1:]
        if cfg:
            model_in = np.concatenate([model_in, model_in], axis=0)
        logits, past = forward_fn(torch.from_numpy(model_in), prm if past is None else None, past)
        scores = logits[:, -1, :].to(torch.float32)
        # transformers' order: default processors (repetition penalty, min length), then the caller's list with the
        # guidance processor last (modeling_parler_tts.py:2556-2567).  Min length is applied after the mix here: the
        # reference's order turns EOS into NaN under guidance (-inf - -inf); see INTEGRATION.md
        if repetition_penalty != 1.0:
            scores = repetition_penalty_(scores, ids, float(repetition_penalty))
        for proc in (logits_processors or []):  # the caller's list runs before the guidance processor appended to it
            scores = proc(torch.from_numpy(ids), scores)
        if cfg:
            scores = classifier_free_guidance(scores, guidance_scale)
        if suppress_from is not None:
            scores = scores.clone()
            scores[:, suppress_from:] = -float("inf")
        if ids.shape[1] < min_cols:
            scores = scores.clone()
            scores[:, eos] = -float("inf")
        if return_trace:
            trace.append(scores.clone())
        tok = select_tokens(scores, do_sample, seed, step, row_offset, temperature, top_k, top_p)
        tok = tok * unfinished + pad * (1 - unfinished)
        ids = np.concatenate([ids, tok[:, None]], axis=1)
        unfinished = unfinished * (tok != eos).astype(np.int64)
        step += 1
        if unfinished.max() == 0 or ids.shape[1] >= max_length:
            break

    codes = dp.revert_delay_pattern(ids, start, pattern, bos, pad, K)
    out = {"sequences": ids, "pattern_mask": pattern, "codes": codes, "steps": step}
    if return_trace:
        out["scores"] = torch.stack(trace, dim=0)  # (steps, R, V)
    return out
