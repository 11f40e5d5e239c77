"""CPU restatement of the T5 text encoder the reference runs once per generate() call
(/root/reference/parler_tts/modeling_parler_tts.py:2229, via AutoModelForTextEncoding -> transformers T5EncoderModel;
the arithmetic is third-party: site-packages/transformers/models/t5/modeling_t5.py).  TEST INFRASTRUCTURE ONLY: imported
by tests/ and oracle/gen_golden.py, never by the product path.  Pinned against transformers' own T5EncoderModel
(tests/golden/t5_tiny.npz, tests/test_oracle_golden.py).

Follows modeling_t5.py: T5LayerNorm 46-68, relative-position buckets 189-234, compute_bias 236-253, T5Attention
254-345 (no 1/sqrt(d) scaling; additive mask finfo.min; fp32 softmax), T5DenseActDense / T5DenseGatedActDense 85-132,
T5LayerSelfAttention / T5LayerFF residual adds, T5Stack final_layer_norm.
"""
import math

import torch


def random_t5_state(cfg, seed=0, std=0.05, ln_jitter=0.1):
    """Seeded HF-key state dict of a T5EncoderModel (weights are never stored in fixtures)."""
    g = torch.Generator().manual_seed(seed)
    D, inner, F = cfg["d_model"], cfg["num_heads"] * cfg["d_kv"], cfg["d_ff"]
    r = lambda *s, sd=std: torch.randn(*s, generator=g) * sd  # noqa: E731
    sd = {"shared.weight": r(cfg["vocab_size"], D, sd=1.0)}
    sd["encoder.embed_tokens.weight"] = sd["shared.weight"]
    for i in range(cfg["num_layers"]):
        p = f"encoder.block.{i}.layer."
        for n in "qkv":
            sd[p + f"0.SelfAttention.{n}.weight"] = r(inner, D)
        sd[p + "0.SelfAttention.o.weight"] = r(D, inner)
        if i == 0:
            sd[p + "0.SelfAttention.relative_attention_bias.weight"] = r(cfg["relative_attention_num_buckets"],
                                                                         cfg["num_heads"], sd=0.5)
        sd[p + "0.layer_norm.weight"] = 1.0 + r(D, sd=ln_jitter)
        if cfg["gated"]:
            sd[p + "1.DenseReluDense.wi_0.weight"] = r(F, D)
            sd[p + "1.DenseReluDense.wi_1.weight"] = r(F, D)
        else:
            sd[p + "1.DenseReluDense.wi.weight"] = r(F, D)
        sd[p + "1.DenseReluDense.wo.weight"] = r(D, F)
        sd[p + "1.layer_norm.weight"] = 1.0 + r(D, sd=ln_jitter)
    sd["encoder.final_layer_norm.weight"] = 1.0 + r(D, sd=ln_jitter)
    return sd


def _rms(x, w, eps):
    var = x.float().pow(2).mean(-1, keepdim=True)
    return w * (x * torch.rsqrt(var + eps))


def _buckets(rel, num_buckets, max_distance):
    nb = num_buckets // 2
    out = (rel > 0).long() * nb
    rp = rel.abs()
    max_exact = nb // 2
    large = max_exact + (torch.log(rp.float() / max_exact) / math.log(max_distance / max_exact)
                         * (nb - max_exact)).long()
    large = torch.min(large, torch.full_like(large, nb - 1))
    return out + torch.where(rp < max_exact, rp, large)


def _act(name):
    if name == "relu":
        return torch.relu
    if name == "gelu_new":
        return lambda x: 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))
    raise ValueError(name)


def t5_encode(sd, cfg, input_ids, attention_mask=None):
    """input_ids (B, S) int64 -> last_hidden_state (B, S, d_model) fp32."""
    B, S = input_ids.shape
    heads, dkv, eps = cfg["num_heads"], cfg["d_kv"], cfg["layer_norm_epsilon"]
    h = sd["shared.weight"].float()[input_ids]
    pos = torch.arange(S)
    rel = pos[None, :] - pos[:, None]  # memory - query
    bias = sd["encoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight"].float()[
        _buckets(rel, cfg["relative_attention_num_buckets"], cfg["relative_attention_max_distance"])]
    bias = bias.permute(2, 0, 1)[None]  # (1, heads, S, S)
    if attention_mask is not None:
        bias = bias + (1.0 - attention_mask[:, None, None, :].float()) * torch.finfo(torch.float32).min
    act = _act(cfg["dense_act_fn"])
    for i in range(cfg["num_layers"]):
        p = f"encoder.block.{i}.layer."
        x = _rms(h, sd[p + "0.layer_norm.weight"].float(), eps)
        split = lambda t: t.view(B, S, heads, dkv).transpose(1, 2)  # noqa: E731
        q, k, v = (split(x @ sd[p + f"0.SelfAttention.{n}.weight"].float().T) for n in "qkv")
        w = torch.softmax(q @ k.transpose(2, 3) + bias, dim=-1)
        a = (w @ v).transpose(1, 2).reshape(B, S, heads * dkv)
        h = h + a @ sd[p + "0.SelfAttention.o.weight"].float().T
        x = _rms(h, sd[p + "1.layer_norm.weight"].float(), eps)
        if cfg["gated"]:
            f = act(x @ sd[p + "1.DenseReluDense.wi_0.weight"].float().T) * (
                x @ sd[p + "1.DenseReluDense.wi_1.weight"].float().T)
        else:
            f = act(x @ sd[p + "1.DenseReluDense.wi.weight"].float().T)
        h = h + f @ sd[p + "1.DenseReluDense.wo.weight"].float().T
    return _rms(h, sd["encoder.final_layer_norm.weight"].float(), eps)
