"""Host side of the CUDA T5 text encoder (ptts_t5_encode): packs a HF `T5EncoderModel` state-dict into the buffers
the C ABI wants and runs the description encoder of `generate()` once per call.

Replaces the third-party torch module the reference calls at /root/reference/parler_tts/modeling_parler_tts.py:2229
(`_prepare_text_encoder_kwargs_for_generation`, 2197-2241).  State-dict keys are HF's (`shared.weight`,
`encoder.block.{i}.layer.0.SelfAttention.{q,k,v,o}.weight`, `encoder.block.0.layer.0.SelfAttention.
relative_attention_bias.weight`, `encoder.block.{i}.layer.{0,1}.layer_norm.weight`,
`encoder.block.{i}.layer.1.DenseReluDense.{wi_0,wi_1,wo | wi,wo}.weight`, `encoder.final_layer_norm.weight`), so a
reference checkpoint's `text_encoder.*` tensors load unchanged.  torch is used for device memory only.
"""
import ctypes as C
import math

import torch

from . import cabi

_ACT = {"relu": cabi.ACT["relu"], "gelu_new": cabi.ACT["gelu_new"], "gelu": cabi.ACT["gelu"], "silu": cabi.ACT["silu"]}
MAX_REL = 512  # table span; every |key - query| >= relative_attention_max_distance shares one bucket per side


def relative_position_buckets(num_buckets, max_distance, max_rel=MAX_REL):
    """Bucket index of every relative position key - query in [-max_rel, max_rel] for a bidirectional stack:
    transformers modeling_t5.py:189-234 evaluated with the same float32 torch ops on the CPU (so boundary
    positions land in the bucket HF puts them in)."""
    rel = torch.arange(-max_rel, max_rel + 1, dtype=torch.long)
    nb = num_buckets // 2
    out = (rel > 0).to(torch.long) * nb
    rp = rel.abs()
    max_exact = nb // 2
    is_small = rp < max_exact
    large = max_exact + (torch.log(rp.float() / max_exact) / math.log(max_distance / max_exact)
                         * (nb - max_exact)).to(torch.long)
    large = torch.min(large, torch.full_like(large, nb - 1))
    return (out + torch.where(is_small, rp, large)).to(torch.int32)


class T5EncoderEngine:
    def __init__(self, t5_config, state_dict, device, dtype):
        cfg = t5_config
        self.device, self.dtype = torch.device(device), dtype
        sd = {k: v.detach() for k, v in state_dict.items()}
        gated = bool(getattr(cfg, "is_gated_act", False))
        act = getattr(cfg, "dense_act_fn", "relu")
        if act not in _ACT:
            raise cabi.ParlerB200Error(f"T5 activation {act!r} is not implemented by the CUDA encoder")
        L, D, heads, dkv, F = cfg.num_layers, cfg.d_model, cfg.num_heads, cfg.d_kv, cfg.d_ff

        def mat(*names):
            return torch.cat([sd[n] for n in names], dim=0).to(self.device, dtype).contiguous()

        def vec(name):
            return sd[name].to(self.device, torch.float32).contiguous()

        emb_key = "shared.weight" if "shared.weight" in sd else "encoder.embed_tokens.weight"
        # parameters stored in bf16 hold bf16-representable values; the residual stream itself is fp32
        self.embed = sd[emb_key].to(self.device, dtype).to(torch.float32).contiguous()
        self.rel_bias = sd["encoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight"].to(
            self.device, dtype).to(torch.float32).contiguous()
        assert self.rel_bias.shape == (cfg.relative_attention_num_buckets, heads), self.rel_bias.shape
        self.bucket_of = relative_position_buckets(cfg.relative_attention_num_buckets,
                                                   cfg.relative_attention_max_distance).to(self.device).contiguous()
        self.ln_f = vec("encoder.final_layer_norm.weight")
        self.layers = []
        arr = (cabi.PttsT5Layer * L)()
        for i in range(L):
            p = f"encoder.block.{i}.layer."
            a = p + "0.SelfAttention."
            t = dict(w_qkv=mat(a + "q.weight", a + "k.weight", a + "v.weight"), w_o=mat(a + "o.weight"),
                     w_wi=(mat(p + "1.DenseReluDense.wi_0.weight", p + "1.DenseReluDense.wi_1.weight") if gated
                           else mat(p + "1.DenseReluDense.wi.weight")),
                     w_wo=mat(p + "1.DenseReluDense.wo.weight"),
                     ln0=vec(p + "0.layer_norm.weight"), ln1=vec(p + "1.layer_norm.weight"))
            self.layers.append(t)
            for name, tensor in t.items():
                setattr(arr[i], name, tensor.data_ptr())
        self._layer_array = arr
        m = cabi.PttsT5()
        m.n_layers, m.d_model, m.n_heads, m.d_kv, m.d_ff = L, D, heads, dkv, F
        m.gated, m.act_fn, m.wdtype = int(gated), _ACT[act], cabi.dtype_enum(dtype)
        m.vocab, m.n_buckets, m.max_rel = self.embed.shape[0], cfg.relative_attention_num_buckets, MAX_REL
        m.eps = float(cfg.layer_norm_epsilon)
        m.embed, m.rel_bias, m.bucket_of = self.embed.data_ptr(), self.rel_bias.data_ptr(), self.bucket_of.data_ptr()
        m.ln_f = self.ln_f.data_ptr()
        m.layers = C.cast(arr, C.POINTER(cabi.PttsT5Layer))
        self.c_model = m
        self.d_model = D
        self._ws = None

    def encode(self, input_ids, attention_mask=None, impl=cabi.IMPL_FAST):
        """input_ids (B, S) int64, attention_mask (B, S) or None -> last_hidden_state (B, S, d_model), compute dtype."""
        if self.device.type != "cuda":
            raise cabi.ParlerB200Error("the T5 encoder runs on the CUDA path only (no CPU fallback)")
        ids = input_ids.to(self.device, torch.int64).contiguous()
        B, S = ids.shape
        mask = None if attention_mask is None else attention_mask.to(self.device, torch.int32).contiguous()
        need = int(cabi.load().ptts_t5_workspace_bytes(C.byref(self.c_model), B * S))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        out = torch.empty(B, S, self.d_model, device=self.device, dtype=self.dtype)
        cabi.call("ptts_t5_encode", C.byref(self.c_model), cabi.ptr(ids), cabi.ptr(mask), B, S, cabi.ptr(self._ws),
                  cabi.ptr(out), impl, cabi.stream_ptr())
        return out
