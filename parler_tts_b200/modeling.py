"""Public model API with the reference's names, arguments and checkpoint layout, driving the sm_100a CUDA path.

Mirrors /root/reference/parler_tts/modeling_parler_tts.py:
  build_delay_pattern_mask 77-139, apply_delay_pattern_mask 68-74, ParlerTTSForCausalLM 1062-1552,
  ParlerTTSForConditionalGeneration 1560-2686, and /root/reference/parler_tts/dac_wrapper/modeling_dac.py (DACModel).
Modules here only HOLD parameters under the reference's state-dict keys; no torch arithmetic runs on the hot path
(the T5 text encoder - a third-party transformers module that runs once per call - is the exception).
"""
import copy
import json
import math
import os
from types import SimpleNamespace

import torch
import torch.nn as nn

from . import cabi
from .configuration import DACConfig, ParlerTTSConfig, ParlerTTSDecoderConfig
from .dac import DacDecoderEngine, DacEncoderEngine
from .engine import DecoderDims, DecoderEngine, DecoderWeights, GenState
from .streaming import AudioChunkStreamer, ChunkPipeline, SilentChunks


# --------------------------------------------------------------------------------------------- helpers
def _cuda_device(t=None):
    if not torch.cuda.is_available():
        raise cabi.ParlerB200Error("parler_tts_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if t is not None and t.is_cuda:
        return t.device
    return torch.device("cuda", torch.cuda.current_device())


def build_delay_pattern_mask(input_ids, bos_token_id, pad_token_id, max_length, num_codebooks):
    """Same contract as the reference function (modeling_parler_tts.py:77-139): returns
    (input_ids[:, :first_start], pattern_mask (B*K, max_length)); computed by ptts_build_delay_mask on the GPU."""
    dev = _cuda_device(input_ids)
    ids = input_ids.reshape(-1, input_ids.shape[-1]).to(dev, torch.int64).contiguous()
    rows, T0 = ids.shape
    B = rows // num_codebooks
    mask = torch.empty(rows, max_length, device=dev, dtype=torch.int64)
    first = torch.zeros(1, device=dev, dtype=torch.int32)
    with torch.cuda.device(dev):
        cabi.call("ptts_build_delay_mask", cabi.ptr(ids), B, num_codebooks, T0, max_length, int(bos_token_id),
                  int(pad_token_id), cabi.ptr(mask), cabi.ptr(first), cabi.stream_ptr())
    if max_length < 2 * num_codebooks - 1:
        out_ids = ids
    else:
        out_ids = mask[:, : int(first.item())]
    return out_ids.to(input_ids.device), mask.to(input_ids.device)


def apply_delay_pattern_mask(input_ids, decoder_pad_token_mask):
    """modeling_parler_tts.py:68-74 via ptts_apply_delay_mask."""
    dev = _cuda_device(input_ids)
    shape = input_ids.shape
    ids = input_ids.reshape(-1, shape[-1]).to(dev, torch.int64).contiguous()
    mask = decoder_pad_token_mask.reshape(-1, decoder_pad_token_mask.shape[-1]).to(dev, torch.int64).contiguous()
    out = torch.empty_like(ids)
    with torch.cuda.device(dev):
        cabi.call("ptts_apply_delay_mask", cabi.ptr(ids), ids.shape[1], cabi.ptr(mask), mask.shape[1], ids.shape[0],
                  ids.shape[1], cabi.ptr(out), out.shape[1], cabi.stream_ptr())
    return out.reshape(shape).to(input_ids.device)


class _Node(nn.Module):
    """Anonymous container so that parameters can live under arbitrary dotted checkpoint keys."""


def _add_param(root, dotted, tensor):
    parts = dotted.split(".")
    mod = root
    for p in parts[:-1]:
        if p not in mod._modules:
            mod.add_module(p, _Node())
        mod = mod._modules[p]
    mod.register_parameter(parts[-1], nn.Parameter(tensor, requires_grad=False))


def sinusoidal_table(num_positions, dim):
    """Position table of the checkpoint (modeling_parler_tts.py:200-213): [cos | sin] halves."""
    half = dim // 2
    freq = torch.exp(torch.arange(half, dtype=torch.int64).float() * -(math.log(10000) / (half - 1)))
    ang = torch.arange(num_positions, dtype=torch.int64).float().unsqueeze(1) * freq.unsqueeze(0)
    tab = torch.cat([torch.cos(ang), torch.sin(ang)], dim=1).view(num_positions, -1)
    if dim % 2 == 1:
        tab = torch.cat([tab, torch.zeros(num_positions, 1)], dim=1)
    return tab


def _load_state_file(path):
    st = os.path.join(path, "model.safetensors")
    if os.path.exists(st):
        from safetensors.torch import load_file

        return load_file(st)
    idx = os.path.join(path, "model.safetensors.index.json")
    if os.path.exists(idx):
        from safetensors.torch import load_file

        out = {}
        for shard in sorted(set(json.load(open(idx))["weight_map"].values())):
            out.update(load_file(os.path.join(path, shard)))
        return out
    pt = os.path.join(path, "pytorch_model.bin")
    if os.path.exists(pt):
        return torch.load(pt, map_location="cpu", weights_only=True)
    raise FileNotFoundError(f"no model.safetensors / pytorch_model.bin under {path}")


def _save_state_file(sd, path):
    from safetensors.torch import save_file

    os.makedirs(path, exist_ok=True)
    out, seen = {}, {}
    for k, v in sd.items():  # tied tensors (T5 `shared` / `encoder.embed_tokens`) are stored once, like HF does
        key = (v.data_ptr(), tuple(v.shape))
        if key in seen and v.numel() > 0:
            continue
        seen[key] = k
        out[k] = v.detach().cpu().contiguous().clone()
    save_file(out, os.path.join(path, "model.safetensors"), metadata={"format": "pt"})


# --------------------------------------------------------------------------------------------- DAC wrapper
class DACModel(nn.Module):
    """Holder of the descript `DAC.state_dict()` under `model.*` (dac_wrapper/modeling_dac.py:13-23) + CUDA decode."""

    config_class = DACConfig

    def __init__(self, config: DACConfig, state_dict=None, decoder_dim=1536, rates=None, codebook_dim=8):
        super().__init__()
        self.config = config
        if rates is None:  # a checkpoint states its own up-sampling rates; random init takes the 44.1 kHz model's
            rates = _dac_rates_of(state_dict) if state_dict is not None else (8, 8, 4, 2)
        self.rates = tuple(int(r) for r in rates)
        self.model = _Node()
        if state_dict is None:
            state_dict = _random_dac_state(config, decoder_dim, self.rates, codebook_dim)
        for k, v in state_dict.items():
            _add_param(self.model, k, v.clone())
        self._engine = None
        self._enc_engine = None

    def engine(self, impl=cabi.IMPL_FAST):
        dev = next(self.parameters()).device
        if self._engine is None or self._engine.device != dev or self._engine.impl != impl:
            self._engine = DacDecoderEngine(dict(self.model.state_dict()), dev, self.rates, impl)
        return self._engine

    def decode(self, audio_codes, audio_scales=None, padding_mask=None, return_dict=None):
        """audio_codes (1, B, K, F) -> audio_values (B, 1, T)   (dac_wrapper/modeling_dac.py:98-134)."""
        if len(audio_codes) != 1:
            raise ValueError(f"Expected one frame, got {len(audio_codes)}")
        audio = self.engine().decode(audio_codes[0])
        if return_dict is False:
            return (audio,)
        return SimpleNamespace(audio_values=audio)

    def decode_chunks(self, audio_codes, chunk_frames=128):
        """Streaming form of `decode` (not in the reference): yields (first_frame,
        audio (B, 1, hop * n)) chunk by chunk; their concatenation equals `decode(audio_codes).audio_values`."""
        if len(audio_codes) != 1:
            raise ValueError(f"Expected one frame, got {len(audio_codes)}")
        yield from self.engine().decode_chunks(audio_codes[0], chunk_frames)

    def encoder_engine(self, impl=cabi.IMPL_FAST):
        dev = next(self.parameters()).device
        if self._enc_engine is None or self._enc_engine.device != dev or self._enc_engine.impl != impl:
            self._enc_engine = DacEncoderEngine(dict(self.model.state_dict()), dev, self.rates, impl)
        return self._enc_engine

    def encode(self, input_values, padding_mask=None, bandwidth=None, return_dict=None, n_quantizers=None,
               sample_rate=None):
        """input_values (B, channels, T) -> audio_codes (1, B, K, ceil(T / hop)), audio_scales [None]
        (dac_wrapper/modeling_dac.py:25-96; one chunk, like the reference's `chunk_length = None`)."""
        if input_values.dim() != 3:
            raise ValueError(f"input_values must be (batch, channels, length), got {tuple(input_values.shape)}")
        channels = input_values.shape[1]
        if channels < 1 or channels > 2:
            raise ValueError(f"Number of audio channels must be 1 or 2, but got {channels}")
        if channels != 1:
            raise ValueError("the DAC encoder takes mono audio (its first conv has one input channel)")
        codes = self.encoder_engine().encode(input_values, n_quantizers)[None]
        if return_dict is False:
            return (codes, [None])
        return SimpleNamespace(audio_codes=codes, audio_scales=[None])

    def forward(self, tensor):
        raise ValueError("`DACModel.forward` not implemented yet")

    # ---- checkpoint I/O as the reference's wrapper is stored (helpers/push_to_hub_scripts/push_dac_to_hub.py:9-13:
    # config.json of DACConfig + the descript `DAC.state_dict()` under `model.*`)
    def save_pretrained(self, path):
        os.makedirs(path, exist_ok=True)
        self.config.save_pretrained(path)
        _save_state_file(self.state_dict(), path)

    @classmethod
    def from_pretrained(cls, path, **kwargs):
        config = DACConfig.from_pretrained(path)
        sd = _load_state_file(path)
        bad = [k for k in sd if not k.startswith("model.")]
        if bad:
            raise RuntimeError(f"not a DACModel checkpoint: keys outside `model.*`, e.g. {bad[:3]}")
        return cls(config, state_dict={k[len("model."):]: v for k, v in sd.items()})


def _dac_rates_of(state_dict):
    """Up-sampling rates of a descript DAC decoder read from its checkpoint: block i holds a transposed conv
    `decoder.model.{i}.block.1` whose kernel is twice its stride (descript DecoderBlock)."""
    rates, i = [], 1
    while True:
        w = state_dict.get(f"decoder.model.{i}.block.1.weight_v", state_dict.get(f"decoder.model.{i}.block.1.weight"))
        if w is None:
            break
        if w.dim() != 3 or w.shape[-1] % 2:
            raise ValueError(f"decoder.model.{i}.block.1: expected a transposed conv with an even kernel, got "
                             f"{tuple(w.shape)}")
        rates.append(w.shape[-1] // 2)
        i += 1
    if not rates:
        raise ValueError("no decoder.model.<i>.block.1 transposed conv in the DAC state dict: cannot infer the rates")
    return tuple(rates)


def _random_dac_state(config, decoder_dim, rates, codebook_dim, seed=0):
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def conv(prefix, cout, cin, k, transposed=False):
        shape = (cin, cout, k) if transposed else (cout, cin, k)
        v = torch.randn(*shape, generator=g) / math.sqrt((cout if transposed else cin) * k)
        sd[prefix + ".weight_v"] = v
        sd[prefix + ".weight_g"] = v.reshape(shape[0], -1).norm(dim=1).reshape(-1, 1, 1)
        sd[prefix + ".bias"] = torch.zeros(cout)

    for i in range(config.num_codebooks):
        q = f"quantizer.quantizers.{i}"
        sd[q + ".codebook.weight"] = torch.randn(config.codebook_size, codebook_dim, generator=g)
        conv(q + ".out_proj", config.latent_dim, codebook_dim, 1)
    ch = decoder_dim
    conv("decoder.model.0", ch, config.latent_dim, 7)
    for bi, s in enumerate(rates):
        p = f"decoder.model.{bi + 1}.block"
        cin, cout = ch // 2 ** bi, ch // 2 ** (bi + 1)
        sd[p + ".0.alpha"] = torch.ones(1, cin, 1)
        conv(p + ".1", cout, cin, 2 * s, transposed=True)
        for ri in range(3):
            r = f"{p}.{ri + 2}.block"
            sd[r + ".0.alpha"] = torch.ones(1, cout, 1)
            conv(r + ".1", cout, cout, 7)
            sd[r + ".2.alpha"] = torch.ones(1, cout, 1)
            conv(r + ".3", cout, cout, 1)
    n = len(rates)
    sd[f"decoder.model.{n + 1}.alpha"] = torch.ones(1, ch // 2 ** n, 1)
    conv(f"decoder.model.{n + 2}", 1, ch // 2 ** n, 7)
    # encoder + in_proj (audio-prompt path): descript Encoder(d_model=64, strides = reversed decoder rates)
    d = max(decoder_dim // 24, 1)
    conv("encoder.block.0", d, 1, 7)
    for bi, s in enumerate(reversed(rates)):
        p = f"encoder.block.{bi + 1}.block"
        for ri in range(3):
            r = f"{p}.{ri}.block"
            sd[r + ".0.alpha"] = torch.ones(1, d, 1)
            conv(r + ".1", d, d, 7)
            sd[r + ".2.alpha"] = torch.ones(1, d, 1)
            conv(r + ".3", d, d, 1)
        sd[p + ".3.alpha"] = torch.ones(1, d, 1)
        conv(p + ".4", 2 * d, d, 2 * s)
        d *= 2
    sd[f"encoder.block.{n + 1}.alpha"] = torch.ones(1, d, 1)
    conv(f"encoder.block.{n + 2}", config.latent_dim, d, 3)
    for i in range(config.num_codebooks):
        conv(f"quantizer.quantizers.{i}.in_proj", codebook_dim, config.latent_dim, 1)
    return sd


# --------------------------------------------------------------------------------------------- generation core
# Defaults of the transformers (<= 4.40) GenerationConfig the reference was written against.  The installed transformers
# leaves unset fields at None; the reference's loop would have seen these values (top_k = 50 is the one that matters:
# `_get_logits_warper`, modeling_parler_tts.py:2597, adds TopKLogitsWarper(50) to every sampling run).
LEGACY_GENERATION_DEFAULTS = dict(
    max_length=20, min_length=0, do_sample=False, num_beams=1, num_beam_groups=1, temperature=1.0, top_k=50, top_p=1.0,
    typical_p=1.0, epsilon_cutoff=0.0, eta_cutoff=0.0, repetition_penalty=1.0, encoder_repetition_penalty=1.0,
    no_repeat_ngram_size=0, num_return_sequences=1, renormalize_logits=False)


def _gcv(gc, name):
    v = getattr(gc, name, None)
    return LEGACY_GENERATION_DEFAULTS.get(name) if v is None else v


def _default_generation_config(dec_cfg):
    from transformers.generation.configuration_utils import GenerationConfig

    return GenerationConfig(bos_token_id=dec_cfg.bos_token_id, pad_token_id=dec_cfg.pad_token_id,
                            eos_token_id=dec_cfg.eos_token_id, decoder_start_token_id=dec_cfg.bos_token_id)


class _nvtx:
    """NVTX range around a phase of generate() (T5 / prefill / loop / DAC): visible in nsys / ncu timelines."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *a):
        torch.cuda.nvtx.range_pop()


def _resolve_generation(self_cfg, generation_config, kwargs, input_len_fn):
    """Steps 1 and 6 of the reference generate (modeling_parler_tts.py:2416-2422, 2499-2527)."""
    gc = copy.deepcopy(generation_config if generation_config is not None else self_cfg)
    model_kwargs = gc.update(**kwargs)
    if gc.pad_token_id is None and gc.eos_token_id is not None:
        eos = gc.eos_token_id
        gc.pad_token_id = eos[0] if isinstance(eos, (list, tuple)) else eos
    if _gcv(gc, "num_beams") != 1 or _gcv(gc, "num_beam_groups") != 1:
        raise ValueError(
            "Got incompatible mode for generation, should be one of greedy or sampling. "
            "Ensure that beam search is de-activated by setting `num_beams=1` and `num_beam_groups=1`.")
    return gc, model_kwargs


def _final_max_length(gc, kwargs, input_len):
    if gc.max_new_tokens is not None:
        gc.max_length = gc.max_new_tokens + input_len
    if gc.max_length is None:
        gc.max_length = LEGACY_GENERATION_DEFAULTS["max_length"]
    if gc.min_length is not None and gc.min_length > gc.max_length:
        raise ValueError(f"Unfeasible length constraints: the minimum length ({gc.min_length}) is larger than"
                         f" the maximum length ({gc.max_length})")
    return int(gc.max_length)


def _first_id(v):
    return v[0] if isinstance(v, (list, tuple)) else v


class _Plan:
    """What the reference's `_get_logits_processor` / `_get_logits_warper` (modeling_parler_tts.py:2561-2567, 2597) would
    build from the generation config, split into what the fused epilogue does natively and what has to run as
    transformers' own processors on the logits tensor between the forward pass and the epilogue (slow loop)."""

    def __init__(self, gc, T0, user_py_procs, cfg, eos_id):
        from transformers.generation import logits_process as lp

        self.do_sample = bool(_gcv(gc, "do_sample"))
        rp = float(_gcv(gc, "repetition_penalty"))
        min_len = int(_gcv(gc, "min_length") or 0)
        if gc.min_new_tokens is not None and gc.min_new_tokens > 0:
            min_len = max(min_len, T0 + int(gc.min_new_tokens))
        if eos_id is None:
            min_len = 0
        erp = float(_gcv(gc, "encoder_repetition_penalty"))
        if erp != 1.0:
            raise NotImplementedError("encoder_repetition_penalty penalises audio-token logits by description token ids; "
                                      "not supported on this path")
        if getattr(gc, "prefix_allowed_tokens_fn", None) is not None or getattr(gc, "constraints", None):
            raise NotImplementedError("constrained decoding is a beam-search feature; this path is greedy / sampling only")
        extra = []  # non-native default processors, in transformers' order
        ngram = int(_gcv(gc, "no_repeat_ngram_size") or 0)
        if ngram > 0:
            extra.append(("ngram", lp.NoRepeatNGramLogitsProcessor(ngram)))
        if getattr(gc, "bad_words_ids", None):
            extra.append(("bad_words", lp.NoBadWordsLogitsProcessor(gc.bad_words_ids, eos_id)))
        if getattr(gc, "forced_bos_token_id", None) is not None:
            extra.append(("forced_bos", lp.ForcedBOSTokenLogitsProcessor(gc.forced_bos_token_id)))
        if getattr(gc, "forced_eos_token_id", None) is not None:
            extra.append(("forced_eos", lp.ForcedEOSTokenLogitsProcessor(int(gc.max_length), gc.forced_eos_token_id, "cuda")))
        if getattr(gc, "exponential_decay_length_penalty", None) is not None:
            extra.append(("exp_decay", lp.ExponentialDecayLengthPenalty(gc.exponential_decay_length_penalty, eos_id, T0)))
        if getattr(gc, "suppress_tokens", None):
            extra.append(("suppress", lp.SuppressTokensLogitsProcessor(gc.suppress_tokens, "cuda")))
        if getattr(gc, "begin_suppress_tokens", None):
            extra.append(("begin_suppress", lp.SuppressTokensAtBeginLogitsProcessor(gc.begin_suppress_tokens, T0, "cuda")))
        if getattr(gc, "sequence_bias", None):
            extra.insert(0, ("sequence_bias", lp.SequenceBiasLogitsProcessor(gc.sequence_bias)))
        renorm = bool(_gcv(gc, "renormalize_logits"))
        self.py_procs = None
        self.rep_penalty, self.min_length = rp, min_len
        if extra or user_py_procs or renorm:
            # everything in python, in the reference's order: defaults, then the caller's list (guidance last, appended by
            # the loop), then the normaliser
            procs = []
            order = ["sequence_bias", "rep", "ngram", "bad_words", "min_len", "forced_bos", "forced_eos", "exp_decay",
                     "suppress", "begin_suppress"]
            named = dict(extra)
            if rp != 1.0:
                named["rep"] = lp.RepetitionPenaltyLogitsProcessor(rp)
            if min_len > 0:
                named["min_len"] = lp.MinLengthLogitsProcessor(min_len, eos_id, "cuda")
            procs = [named[n] for n in order if n in named] + list(user_py_procs)
            self.py_procs, self.py_renorm = procs, renorm
            self.rep_penalty, self.min_length = 1.0, 0
        # warpers (sampling only): temperature, top-k, top-p natively; the rest through transformers' classes
        self.temperature, self.top_k, self.top_p, self.py_warpers = 1.0, 0, 1.0, None
        if self.do_sample:
            temp, top_k, top_p = float(_gcv(gc, "temperature")), int(_gcv(gc, "top_k") or 0), float(_gcv(gc, "top_p"))
            if temp <= 0:
                raise ValueError(f"`temperature` (={temp}) has to be a strictly positive float")
            if not 0 < top_p <= 1.0:
                raise ValueError(f"`top_p` has to be a float > 0 and <= 1, but is {top_p}")
            warp = []
            min_p = getattr(gc, "min_p", None)
            if min_p is not None:
                warp.append(lp.MinPLogitsWarper(float(min_p)))
            if float(_gcv(gc, "typical_p")) < 1.0:
                warp.append(lp.TypicalLogitsWarper(mass=float(gc.typical_p)))
            if float(_gcv(gc, "epsilon_cutoff")) > 0.0:
                warp.append(lp.EpsilonLogitsWarper(float(gc.epsilon_cutoff)))
            if float(_gcv(gc, "eta_cutoff")) > 0.0:
                warp.append(lp.EtaLogitsWarper(float(gc.eta_cutoff), device="cuda"))
            if warp or renorm:
                head = []
                if temp != 1.0:
                    head.append(lp.TemperatureLogitsWarper(temp))
                if top_k > 0:
                    head.append(lp.TopKLogitsWarper(top_k))
                if top_p < 1.0:
                    head.append(lp.TopPLogitsWarper(top_p))
                self.py_warpers = head + warp
                if self.py_procs is None:
                    self.py_procs, self.py_renorm = [], renorm
            else:
                self.temperature, self.top_k, self.top_p = temp, top_k, top_p
        self.cfg_in_python = cfg and self.py_procs is not None

    @property
    def slow(self):
        return self.py_procs is not None


def _draw_seed():
    # follows torch's global generator so that `set_seed(s); model.generate(do_sample=True)` is reproducible
    return int(torch.randint(0, 2 ** 62, (1,)).item())


def _special_processors(logits_processor):
    """Splits a user LogitsProcessorList into (suppress_from handled in-kernel, remaining python callables)."""
    suppress, rest = -1, []
    for p in (logits_processor or []):
        if isinstance(p, SuppressTokensAtOrAbove):
            suppress = p.first_banned if suppress < 0 else min(suppress, p.first_banned)
        else:
            rest.append(p)
    return suppress, rest


class SuppressTokensAtOrAbove:
    """A logits processor the fused epilogue understands natively: scores[:, first_banned:] = -inf.
    (Keeps random-init models away from EOS/BOS/unused ids)  Also callable like any HF
    processor."""

    def __init__(self, first_banned):
        self.first_banned = int(first_banned)

    def __call__(self, input_ids, scores):
        scores = scores.clone()
        scores[:, self.first_banned:] = -float("inf")
        return scores


def _drop_if_all_ones(mask):
    """An attention mask without a zero masks nothing: None selects the kernels without the mask loads (the same sums
    in the same order).  One small reduction + host read; callers do it before they queue long work."""
    return None if mask is None or bool(mask.all()) else mask


def _generate_codes(decoder, enc, enc_mask, prompt_table, prompt_ids, prompt_hidden, prompt_mask, decoder_input_ids,
                    gc, kwargs, logits_processor, stopping_criteria, streamer, impl, use_graph, seed, row_offset,
                    audio_engine=None):
    """The loop of transformers greedy_search / sample on the CUDA path.  Returns (codes (Bu,K,F), sequences, state).
    With an `AudioChunkStreamer` and the codec engine (`audio_engine`), finished frames are de-delayed and decoded on
    a second stream while the loop runs (streaming.ChunkPipeline, left on `state.audio_pipeline` for the caller)."""
    w = decoder.packed_weights()
    d = w.dims
    K = d.num_codebooks
    dev = w.device
    cfg = gc.guidance_scale is not None and gc.guidance_scale > 1
    B = enc.shape[0] if enc is not None else (prompt_ids if prompt_ids is not None else prompt_hidden).shape[0]
    Bu = B // 2 if cfg else B
    bos = gc.decoder_start_token_id if gc.decoder_start_token_id is not None else gc.bos_token_id
    if bos is None:
        bos = d.bos_token_id
    pad = gc.pad_token_id if gc.pad_token_id is not None else d.pad_token_id
    eos = _first_id(gc.eos_token_id)  # None: rows never finish (the loop runs to max_length), as in the reference
    if isinstance(gc.eos_token_id, (list, tuple)) and len(gc.eos_token_id) > 1:
        raise NotImplementedError("a list of several eos_token_id values is not supported by the fused epilogue")
    start = torch.full((Bu * K, 1), bos, dtype=torch.int64, device=dev)
    if decoder_input_ids is not None:
        user = decoder_input_ids.to(dev, torch.int64).reshape(Bu * K, -1)
        if bool((user[..., 0] != bos).all().item()):  # modeling_parler_tts.py:2185-2186
            user = torch.cat([start, user], dim=-1)
        start = user
    T0 = start.shape[1]
    max_length = _final_max_length(gc, kwargs, T0)
    if max_length < 2 * K - 1:
        raise ValueError(f"max_length={max_length} is shorter than 2*num_codebooks-1: the delay pattern cannot be "
                         "applied (modeling_parler_tts.py:105-106) and no frame would survive the de-delay")
    if T0 >= max_length:
        raise ValueError(f"Input length of decoder_input_ids is {T0}, but `max_length` is set to {max_length}")
    P = 0 if (prompt_ids is None and prompt_hidden is None) else (prompt_ids if prompt_ids is not None
                                                                    else prompt_hidden).shape[1]
    if P + max_length > d.max_position_embeddings:
        raise ValueError(f"prompt length {P} + max_length {max_length} exceeds max_position_embeddings "
                         f"{d.max_position_embeddings}")
    suppress, user_py = _special_processors(logits_processor)
    plan = _Plan(gc, T0, user_py, cfg, eos)
    do_sample, top_k, temperature = plan.do_sample, plan.top_k, plan.temperature
    S = 0 if enc is None else enc.shape[1]
    gscale = float(gc.guidance_scale) if cfg else 1.0
    kv_dtype = getattr(decoder, "kv_dtype", None)
    sak = int(getattr(decoder, "self_attn_kernel", 0) or 0)
    key = (sak, id(w), B, S, P, T0, max_length, impl, do_sample, top_k, temperature, suppress, gscale, row_offset,
           enc_mask is not None, prompt_mask is not None, kv_dtype, plan.min_length, plan.top_p, plan.rep_penalty,
           plan.cfg_in_python, bos, pad, eos)
    cached = getattr(decoder, "_state_cache", None)
    if cached is not None and cached[0] == key:  # same shapes: reuse buffers and the captured CUDA graph
        st = cached[1]
        if enc_mask is not None:
            st.enc_mask.copy_(enc_mask.to(dev, torch.int32))
        if prompt_mask is not None:
            st.prompt_mask.copy_(prompt_mask.to(dev, torch.int32))
    else:
        decoder._state_cache = None
        st = GenState(w, B, S, P, T0, max_length, impl=impl, do_sample=do_sample, top_k=top_k,
                      temperature=temperature, suppress_from=suppress, guidance_scale=gscale, seed=seed,
                      row_offset=row_offset * K,
                      enc_mask=None if enc_mask is None else enc_mask.to(dev, torch.int32).clone(),
                      prompt_mask=None if prompt_mask is None else prompt_mask.to(dev, torch.int32).clone(),
                      kv_dtype=kv_dtype, self_attn_kernel=sak, min_length=plan.min_length, top_p=plan.top_p,
                      repetition_penalty=plan.rep_penalty, pad_token_id=pad,
                      eos_token_id=-1 if eos is None else eos, bos_token_id=bos)
        st.c_state.logits_premixed = int(plan.cfg_in_python)
        decoder._state_cache = (key, st)
    eng = DecoderEngine(w)
    st.reset(start, seed)
    audio_streamer = streamer if isinstance(streamer, AudioChunkStreamer) else None
    if audio_streamer is not None:
        streamer = None  # audio chunks, not token ids: the token-level hook of the host loop is not needed
    if streamer is not None:
        streamer.put(start.cpu())
    if enc is not None:
        eng.cross_kv(st, enc)
    slow = plan.slow or streamer is not None or bool(stopping_criteria)
    st.audio_pipeline = None
    if not slow:
        if audio_streamer is not None and audio_engine is not None and T0 == 1:
            st.audio_pipeline = ChunkPipeline(audio_engine, st, audio_streamer, K)
        with _nvtx("ptts.prefill"):
            eng.prefill(st, prompt_table, prompt_ids, prompt_hidden)
        with _nvtx("ptts.decode_loop"):
            ctrl = eng.run_loop(st, use_graph=use_graph,
                                on_progress=st.audio_pipeline.progress if st.audio_pipeline is not None else None)
    else:
        with _nvtx("ptts.decode_loop_host_processors"):
            ctrl = _slow_loop(eng, st, prompt_table, prompt_ids, prompt_hidden, plan, gscale, stopping_criteria,
                              streamer)
    Lout = ctrl["stop_len"] if ctrl["done"] else ctrl["len"]
    seq = st.ids[:, :Lout]
    if T0 == 1:
        codes = torch.empty(Bu, K, Lout if Lout < 2 * K - 1 else Lout - K, device=dev, dtype=torch.int64)
        cabi.call("ptts_dedelay", cabi.ptr(st.ids), st.max_len, cabi.ptr(st.mask), st.max_len, Bu, K, Lout,
                  cabi.ptr(codes), cabi.stream_ptr())
    else:  # audio-prompt geometry: the general boolean-mask form of modeling_parler_tts.py:2637-2645
        applied = apply_delay_pattern_mask(seq, st.mask)
        _, m2 = build_delay_pattern_mask(start, bos, pad, Lout, K)
        keep = (m2 != bos) & (m2 != pad)
        codes = applied[keep].reshape(Bu, K, -1)
    return codes, seq, st


def _slow_loop(eng, st, prompt_table, prompt_ids, prompt_hidden, plan, gscale, stopping_criteria, streamer):
    """Step-at-a-time variant used when host-side logits processors / warpers, stopping criteria or a streamer are
    supplied: forward (kernels) -> transformers' processors on the logits tensor, in the reference's order -> fused
    epilogue kernel (selection, EOS / PAD bookkeeping, delay mask, next embedding)."""
    procs = plan.py_procs or []
    warpers = plan.py_warpers or []
    first = True
    rows = st.ids.shape[0]  # Bu * K
    while True:
        if first:
            eng.prefill(st, prompt_table, prompt_ids, prompt_hidden, run_epilogue=False)
            first = False
        else:
            eng.step(st, run_epilogue=False)
        c = st.read_ctrl()
        hist = st.ids[:, : c["len"]]
        if plan.slow:
            scores = st.logits
            for p in procs:  # with guidance the scores hold both halves, the ids one (as in the reference)
                scores = p(hist, scores)
            if plan.cfg_in_python:  # ClassifierFreeGuidanceLogitsProcessor, appended last (modeling_parler_tts.py:2556-2558)
                lsm = torch.log_softmax(scores.float(), dim=-1)
                cond, unc = lsm[:rows], lsm[rows:]
                scores = unc + (cond - unc) * gscale
            if plan.py_renorm:
                scores = torch.log_softmax(scores.float(), dim=-1)
            for wproc in warpers:
                scores = wproc(hist, scores)
            if warpers and plan.py_renorm:
                scores = torch.log_softmax(scores.float(), dim=-1)
            st.logits[: scores.shape[0]].copy_(scores)
        eng.epilogue(st)
        c = st.read_ctrl()
        hist = st.ids[:, : c["len"]]
        if streamer is not None:
            streamer.put(hist[:, -1].cpu())
        stop = bool(c["done"])
        for crit in (stopping_criteria or []):
            r = crit(hist, st.logits)
            stop = stop or bool(r.all() if hasattr(r, "all") else r)
        if stop:
            if streamer is not None:
                streamer.end()
            c["done"], c["stop_len"] = 1, c["len"]
            return c


# --------------------------------------------------------------------------------------------- decoder-only model
class ParlerTTSForCausalLM(nn.Module):
    """Decoder + 9 lm_heads (modeling_parler_tts.py:1062-1552).  State-dict keys: `model.decoder.*`, `lm_heads.{k}.*`."""

    config_class = ParlerTTSDecoderConfig

    def __init__(self, config: ParlerTTSDecoderConfig, state_dict=None, seed=0):
        super().__init__()
        self.config = config
        self.num_codebooks = config.num_codebooks
        self.generation_config = _default_generation_config(config)
        dims = DecoderDims.from_config(config)
        if state_dict is None:
            state_dict = _random_decoder_state(dims, config.initializer_factor, seed)
        self.model, self.lm_heads = _Node(), _Node()
        for k, v in state_dict.items():
            if k.startswith("model."):
                _add_param(self.model, k[len("model."):], v.clone())
            elif k.startswith("lm_heads."):
                _add_param(self.lm_heads, k[len("lm_heads."):], v.clone())
        self._packed = None
        self._state_cache = None
        self.impl = cabi.IMPL_FAST
        self.use_cuda_graph = True
        # self-attention KV cache element type: None = the weight dtype; "fp8" = e4m3 with per-(layer, K|V, head) scales
        # (bf16 weights + fast kernels only; halves the bytes of the dominant stream)
        self.kv_dtype = None
        # decode-step self-attention kernel (PttsState.self_attn_kernel): 0 automatic, 1 one CTA per (utterance, head),
        # 2 the persistent warp-specialised tcgen05 kernel whenever its shape constraints hold
        self.self_attn_kernel = 0

    # ---- plumbing
    @property
    def device(self):
        return next(self.parameters()).device

    @property
    def dtype(self):
        return self.lm_heads._modules["0"].weight.dtype

    def packed_weights(self):
        dev, dt = self.device, self.dtype
        if dev.type != "cuda":
            raise cabi.ParlerB200Error("move the model to a CUDA device first: there is no CPU path")
        if self._packed is None or self._packed.device != dev or self._packed.dtype != dt:
            sd = {k: v for k, v in self.state_dict().items()}
            self._packed = DecoderWeights(sd, DecoderDims.from_config(self.config), dev, dt)
        return self._packed

    def reset_engine(self):
        self._packed, self._state_cache = None, None

    # ---- checkpoint I/O of the stand-alone decoder (config.json of ParlerTTSDecoderConfig + `model.decoder.*` /
    # `lm_heads.*`), the form helpers/model_init_scripts/init_model_600M.py:44-45 saves before it assembles the composite
    def save_pretrained(self, path):
        os.makedirs(path, exist_ok=True)
        self.config.save_pretrained(path)
        _save_state_file(self.state_dict(), path)

    @classmethod
    def from_pretrained(cls, path, torch_dtype=None, config=None, **kwargs):
        if config is None:
            config = ParlerTTSDecoderConfig.from_pretrained(path)
        if isinstance(config, ParlerTTSConfig):  # a composite checkpoint: take its decoder (modeling_parler_tts.py:1893-1894)
            config = config.decoder
        sd = _load_state_file(path)
        if any(k.startswith("decoder.") for k in sd):
            sd = {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}
        bad = [k for k in sd if not k.startswith(("model.", "lm_heads."))]
        if bad:
            raise RuntimeError(f"not a ParlerTTSForCausalLM checkpoint: unexpected keys, e.g. {bad[:3]}")
        model = cls(config, state_dict=sd)
        return model if torch_dtype is None else model.to(torch_dtype)

    def build_delay_pattern_mask(self, input_ids, bos_token_id, pad_token_id, max_length=None):
        max_length = max_length if max_length is not None else self.generation_config.max_length
        return build_delay_pattern_mask(input_ids, bos_token_id, pad_token_id, max_length, self.num_codebooks)

    @staticmethod
    def apply_delay_pattern_mask(input_ids, decoder_pad_token_mask):
        return apply_delay_pattern_mask(input_ids, decoder_pad_token_mask)

    # ---- forward: teacher-forced logits, optionally with the static cache handed back as `past_key_values`
    @torch.no_grad()
    def forward(self, input_ids=None, attention_mask=None, encoder_hidden_states=None, encoder_attention_mask=None,
                prompt_hidden_states=None, prompt_attention_mask=None, past_key_values=None, use_cache=None,
                max_cache_len=None, **unused):
        """logits (B*K, T, vocab) as ParlerTTSForCausalLM.forward without labels (modeling_parler_tts.py:1096-1185).

        First call (past_key_values=None): prompt + all given columns in one pass, logits of every position
        (B*K, P+T, vocab).  With use_cache=True the static cache comes back as `past_key_values` (a `StaticDecoderCache`,
        not the reference's tuple of per-layer tensors) and a later call
        `forward(input_ids=next_column(s), past_key_values=cache)` runs one cached decode step per column through the
        same kernels generate() uses (encoder / prompt arguments are then ignored, as the reference's cached branch
        ignores them, modeling_parler_tts.py:292-294, 2132-2133)."""
        if attention_mask is not None:
            raise NotImplementedError("decoder attention_mask is not supported; pass prompt_attention_mask")
        w = self.packed_weights()
        K = self.num_codebooks
        ids = input_ids.to(w.device, torch.int64).reshape(-1, input_ids.shape[-1]).contiguous()
        B, T = ids.shape[0] // K, ids.shape[1]
        eng = DecoderEngine(w)
        if past_key_values is not None:
            if not isinstance(past_key_values, StaticDecoderCache):
                raise TypeError("past_key_values must be the StaticDecoderCache an earlier forward(use_cache=True) "
                                "returned; tuples of per-layer tensors are not supported (INTEGRATION.md)")
            st = past_key_values.state
            if B != st.B:
                raise ValueError(f"batch {B} does not match the cache's batch {st.B}")
            V = self.config.vocab_size
            out = torch.empty(B * K, T, V, device=w.device, dtype=torch.float32)
            for t in range(T):
                n = past_key_values.columns
                if n + 1 > st.max_len:
                    raise ValueError(f"cache is full ({st.max_len} columns); pass a larger max_cache_len")
                st.ids[:, n] = ids[:, t]
                st.ctrl[0] = n + 1
                cabi.call("ptts_embed_tokens", eng._m(), cabi.ptr(st.ids[:, n:]), st.max_len, B, 1, st.P + n,
                          cabi.ptr(st.h), cabi.stream_ptr())
                eng.step(st, run_epilogue=False)
                out[:, t] = st.logits
                past_key_values.columns = n + 1
            return SimpleNamespace(logits=out, past_key_values=past_key_values)
        P = 0 if prompt_hidden_states is None else prompt_hidden_states.shape[1]
        S = 0 if encoder_hidden_states is None else encoder_hidden_states.shape[1]
        cap = max(T + 1, 2)
        if use_cache:
            cap = int(max_cache_len) if max_cache_len is not None else min(
                self.config.max_position_embeddings - P, T + 512)
            if cap < T + 1:
                raise ValueError(f"max_cache_len={cap} is too small for {T} columns")
        st = GenState(w, B, S, P, T, cap, impl=self.impl, enc_mask=encoder_attention_mask,
                      prompt_mask=prompt_attention_mask, kv_dtype=self.kv_dtype if use_cache else None,
                      self_attn_kernel=self.self_attn_kernel)
        st.mask.fill_(-1)
        st.ids[:, :T] = ids
        st.ctrl.copy_(torch.tensor([T, 0, 0, B * K, 0, 0, 0, 0], dtype=torch.int32))
        if S:
            eng.cross_kv(st, encoder_hidden_states)
        ph = None if prompt_hidden_states is None else prompt_hidden_states.to(w.device, torch.float32).contiguous()
        logits = eng.prefill(st, prompt_hidden=ph, all_logits=True, run_epilogue=False)
        cache = StaticDecoderCache(st, T) if use_cache else None
        return SimpleNamespace(logits=logits, past_key_values=cache)

    @torch.no_grad()
    def generate(self, inputs=None, generation_config=None, logits_processor=None, stopping_criteria=None,
                 synced_gpus=None, streamer=None, **kwargs):
        """Decoder-only generation -> codes (B, K, F)   (modeling_parler_tts.py:1281-1552)."""
        gc, mk = _resolve_generation(self.generation_config, generation_config, kwargs, None)
        seed = mk.pop("seed", None)
        enc = mk.pop("encoder_hidden_states", None)
        enc_mask = mk.pop("encoder_attention_mask", None)
        prm = mk.pop("prompt_hidden_states", None)
        prm_mask = mk.pop("prompt_attention_mask", None)
        dec_ids = inputs if inputs is not None else mk.pop("input_ids", None)
        w = self.packed_weights()
        if enc is None and prm is None and dec_ids is None:
            raise ValueError("generate() needs encoder_hidden_states, prompt_hidden_states or input_ids")
        cfg = gc.guidance_scale is not None and gc.guidance_scale > 1
        if cfg:  # modeling_parler_tts.py:1214-1229: zero prompt for the unconditional half
            if prm is not None:
                prm = torch.cat([prm, torch.zeros_like(prm)], dim=0)
            if prm_mask is not None:
                prm_mask = torch.cat([prm_mask, torch.zeros_like(prm_mask)], dim=0)
            # encoder_hidden_states / encoder_attention_mask must already hold both halves (as in the reference,
            # where the composite model doubles them in _prepare_text_encoder_kwargs_for_generation)
        enc_mask, prm_mask = _drop_if_all_ones(enc_mask), _drop_if_all_ones(prm_mask)
        if prm is not None:
            prm = prm.to(w.device, torch.float32).contiguous()
        if enc is None and prm is None:
            B = dec_ids.shape[0] // self.num_codebooks
            prm = torch.zeros(B, 0, self.config.hidden_size, device=w.device)
        codes, seq, _ = _generate_codes(self, enc, enc_mask, None, None, prm, prm_mask, dec_ids, gc, kwargs,
                                        logits_processor, stopping_criteria, streamer, self.impl, self.use_cuda_graph,
                                        _draw_seed() if seed is None else seed, 0)
        if gc.return_dict_in_generate:
            return SimpleNamespace(sequences=codes)
        return codes


class StaticDecoderCache:
    """What ParlerTTSForCausalLM.forward(use_cache=True) returns as `past_key_values`: the preallocated self-attention
    cache [L][2][B][heads][P + max_cache_len][d], the cross-attention cache and the token matrix of one batch (a
    GenState), plus the number of token columns already in it.  Replaces the reference's growing tuple of per-layer
    (self_k, self_v, cross_k, cross_v) tensors (modeling_parler_tts.py:303-304, 501)."""

    def __init__(self, state, columns):
        self.state, self.columns = state, int(columns)

    def get_seq_length(self):
        return self.state.P + self.columns


def _random_decoder_state(dims, std, seed):
    """Random init as the reference's `_init_weights` (modeling_parler_tts.py:517-526): N(0, std), LayerNorm 1/0."""
    g = torch.Generator().manual_seed(seed)
    H, Fd, V, K = dims.hidden_size, dims.ffn_dim, dims.vocab_size, dims.num_codebooks
    sd = {}
    nrm = lambda *s: torch.randn(*s, generator=g) * std  # noqa: E731
    for k in range(K):
        sd[f"model.decoder.embed_tokens.{k}.weight"] = nrm(V + 1, H)
    sd["model.decoder.embed_positions.weights"] = sinusoidal_table(dims.max_position_embeddings, H)
    for i in range(dims.num_hidden_layers):
        p = f"model.decoder.layers.{i}."
        for att in ("self_attn", "encoder_attn"):
            for proj in ("q_proj", "k_proj", "v_proj", "out_proj"):
                sd[p + f"{att}.{proj}.weight"] = nrm(H, H)
        for ln in ("self_attn_layer_norm", "encoder_attn_layer_norm", "final_layer_norm"):
            sd[p + ln + ".weight"], sd[p + ln + ".bias"] = torch.ones(H), torch.zeros(H)
        sd[p + "fc1.weight"], sd[p + "fc2.weight"] = nrm(Fd, H), nrm(H, Fd)
    sd["model.decoder.layer_norm.weight"], sd["model.decoder.layer_norm.bias"] = torch.ones(H), torch.zeros(H)
    for k in range(K):
        sd[f"lm_heads.{k}.weight"] = nrm(V, H)
    return sd


# --------------------------------------------------------------------------------------------- composite model
class ParlerTTSForConditionalGeneration(nn.Module):
    """text encoder (T5) + DAC + decoder (modeling_parler_tts.py:1560-2686).  State-dict keys: `text_encoder.*`,
    `audio_encoder.model.*`, `decoder.*`, `enc_to_dec_proj.{weight,bias}`, `embed_prompts.weight`."""

    config_class = ParlerTTSConfig
    main_input_name = "input_ids"

    def __init__(self, config=None, text_encoder=None, audio_encoder=None, decoder=None, seed=0):
        super().__init__()
        if config is None and (text_encoder is None or audio_encoder is None or decoder is None):
            raise ValueError("Either a configuration has to be provided, or all three of text encoder, audio encoder "
                             "and Parler-TTS decoder.")
        if config is None:
            config = ParlerTTSConfig.from_sub_models_config(text_encoder.config, audio_encoder.config, decoder.config)
        elif not isinstance(config, ParlerTTSConfig):
            raise ValueError(f"Config: {config} has to be of type {ParlerTTSConfig}")
        self.config = config
        if text_encoder is None:
            from transformers import AutoModelForTextEncoding

            text_encoder = AutoModelForTextEncoding.from_config(config.text_encoder)
        self.text_encoder = text_encoder.eval()
        self.audio_encoder = audio_encoder if audio_encoder is not None else DACModel(config.audio_encoder)
        self.decoder = decoder if decoder is not None else ParlerTTSForCausalLM(config.decoder, seed=seed)
        g = torch.Generator().manual_seed(seed + 1)
        std = config.decoder.initializer_factor
        H, He = config.decoder.hidden_size, config.text_encoder.hidden_size
        if He != H:
            self.enc_to_dec_proj = _Node()
            _add_param(self.enc_to_dec_proj, "weight", torch.randn(H, He, generator=g) * std)
            _add_param(self.enc_to_dec_proj, "bias", torch.zeros(H))
        self.embed_prompts = _Node()
        _add_param(self.embed_prompts, "weight", torch.randn(config.vocab_size, H, generator=g) * std)
        self.generation_config = _default_generation_config(config.decoder)
        self._prompt_table = None
        self.text_encoder_impl = "native"  # CUDA T5 encoder (ptts_t5_encode); "torch" = the third-party module
        # codec decode beside the token loop when no streamer was passed (streaming.SilentChunks): frames per chunk
        # (0 = decode after the loop) and the CTAs a chunk decode may occupy while the loop runs (None = one per SM)
        self.overlap_decode_chunk_frames = 0
        self.overlap_decode_ctas = None
        self._t5_engine = None

    # ---- reference accessors
    def get_audio_encoder(self):
        return self.audio_encoder

    def get_text_encoder(self):
        return self.text_encoder

    def get_encoder(self):
        return self.text_encoder

    def get_decoder(self):
        return self.decoder

    @property
    def device(self):
        return self.decoder.device

    @property
    def dtype(self):
        return self.decoder.dtype

    # ---- checkpoint I/O (config.json + generation_config.json + model.safetensors, reference key names)
    @classmethod
    def from_pretrained(cls, path, torch_dtype=None, **kwargs):
        from transformers.generation.configuration_utils import GenerationConfig

        config = ParlerTTSConfig.from_pretrained(path)
        sd = _load_state_file(path)
        dec_sd = {k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}
        dac_sd = {k[len("audio_encoder.model."):]: v for k, v in sd.items() if k.startswith("audio_encoder.model.")}
        model = cls(config, decoder=ParlerTTSForCausalLM(config.decoder, state_dict=dec_sd),
                    audio_encoder=DACModel(config.audio_encoder, state_dict=dac_sd))
        rest = {k: v for k, v in sd.items() if not k.startswith(("decoder.", "audio_encoder."))}
        missing, unexpected = model.load_state_dict(rest, strict=False)
        missing = [k for k in missing if not k.startswith(("decoder.", "audio_encoder."))
                   and not k.endswith("encoder.embed_tokens.weight")]  # tied to text_encoder.shared.weight
        if missing or unexpected:
            raise RuntimeError(f"checkpoint mismatch: missing {missing[:5]} unexpected {unexpected[:5]}")
        if os.path.exists(os.path.join(path, "generation_config.json")):
            model.generation_config = GenerationConfig.from_pretrained(path)
        if torch_dtype is not None:
            model = model.to(torch_dtype)
        return model

    def save_pretrained(self, path):
        os.makedirs(path, exist_ok=True)
        self.config.save_pretrained(path)
        self.generation_config.save_pretrained(path)
        _save_state_file(self.state_dict(), path)

    @classmethod
    def from_sub_models_pretrained(cls, text_encoder_pretrained_model_name_or_path=None,
                                   audio_encoder_pretrained_model_name_or_path=None,
                                   decoder_pretrained_model_name_or_path=None, *model_args, **kwargs):
        """modeling_parler_tts.py:1722-1925: a text encoder, a DAC wrapper and a decoder from their own checkpoints (or
        passed as `text_encoder_model=` / `audio_encoder_model=` / `decoder_model=`), `enc_to_dec_proj` and
        `embed_prompts` freshly initialised; remaining kwargs (`vocab_size=...`) go to the composite config."""
        def split(prefix):
            own = {k[len(prefix):]: v for k, v in kwargs.items() if k.startswith(prefix)}
            for k in own:
                del kwargs[prefix + k]
            return own

        kw_text, kw_audio, kw_dec = split("text_encoder_"), split("audio_encoder_"), split("decoder_")
        text_encoder = kw_text.pop("model", None)
        if text_encoder is None:
            if text_encoder_pretrained_model_name_or_path is None:
                raise ValueError("If `text_encoder_model` is not defined as an argument, a "
                                 "`text_encoder_pretrained_model_name_or_path` has to be defined.")
            from transformers import AutoModelForTextEncoding

            text_encoder = AutoModelForTextEncoding.from_pretrained(text_encoder_pretrained_model_name_or_path,
                                                                    *model_args, **kw_text)
        audio_encoder = kw_audio.pop("model", None)
        if audio_encoder is None:
            if audio_encoder_pretrained_model_name_or_path is None:
                raise ValueError("If `audio_encoder_model` is not defined as an argument, an "
                                 "`audio_encoder_pretrained_model_name_or_path` has to be defined.")
            audio_encoder = DACModel.from_pretrained(audio_encoder_pretrained_model_name_or_path, **kw_audio)
        decoder = kw_dec.pop("model", None)
        if decoder is None:
            if decoder_pretrained_model_name_or_path is None:
                raise ValueError("If `decoder_model` is not defined as an argument, a "
                                 "`decoder_pretrained_model_name_or_path` has to be defined.")
            decoder = ParlerTTSForCausalLM.from_pretrained(decoder_pretrained_model_name_or_path, **kw_dec)
        config = ParlerTTSConfig.from_sub_models_config(text_encoder.config, audio_encoder.config, decoder.config, **kwargs)
        return cls(text_encoder=text_encoder, audio_encoder=audio_encoder, decoder=decoder, config=config)

    def prompt_table(self):
        t = self.embed_prompts.weight
        if self._prompt_table is None or self._prompt_table.device != t.device:
            self._prompt_table = t.detach().to(torch.float32).contiguous()
        return self._prompt_table

    # ---- pieces of generate
    @torch.no_grad()
    def encode_text(self, input_ids, attention_mask):
        """T5 description encoder, once per call (modeling_parler_tts.py:2229): the CUDA encoder for a T5 text_encoder,
        else the third-party torch module.  enc_to_dec_proj + bias + attention-mask multiply follow in
        project_encoder (modeling_parler_tts.py:2010-2017, hoisted out of the per-step loop)."""
        eng = self.text_encoder_engine()
        if eng is not None:
            return eng.encode(input_ids, attention_mask, impl=self.decoder.impl)
        hs = self.text_encoder(input_ids=input_ids, attention_mask=attention_mask, return_dict=True).last_hidden_state
        return hs

    def text_encoder_engine(self):
        """CUDA T5 encoder over the text_encoder module's own parameters (HF key names), rebuilt when they move or
        change dtype.  `text_encoder_impl = "torch"` keeps the third-party module (any non-T5 encoder does too)."""
        te = self.text_encoder
        if self.text_encoder_impl != "native" or getattr(te.config, "model_type", None) != "t5":
            return None
        p = next(te.parameters())
        if p.device.type != "cuda" or p.dtype not in (torch.float32, torch.bfloat16):
            return None
        key = (p.device, p.dtype, p.data_ptr())
        if self._t5_engine is None or self._t5_engine[0] != key:
            from .text_encoder import T5EncoderEngine

            self._t5_engine = (key, T5EncoderEngine(te.config, te.state_dict(), p.device, p.dtype))
        return self._t5_engine[1]

    def project_encoder(self, hs, attention_mask):
        w = self.decoder.packed_weights()
        B, S, He = hs.shape
        H = self.config.decoder.hidden_size
        dt = w.dtype
        rm = None if attention_mask is None else attention_mask.to(w.device, torch.int32).reshape(-1).contiguous()
        if hasattr(self, "enc_to_dec_proj"):
            x = hs.to(w.device, dt).contiguous()
            wt = self.enc_to_dec_proj.weight.detach().to(w.device, dt).contiguous()
            bias = self.enc_to_dec_proj.bias.detach().to(w.device, torch.float32).contiguous()
            y = torch.empty(B, S, H, device=w.device, dtype=dt)
            cabi.call("ptts_linear", cabi.ptr(x), cabi.dtype_enum(dt), cabi.ptr(wt), cabi.dtype_enum(dt),
                      cabi.ptr(bias), B * S, H, He, 1.0, 0, None, cabi.ptr(rm), cabi.ptr(y), cabi.dtype_enum(dt),
                      self.decoder.impl, cabi.stream_ptr())
            return y
        y = hs.to(w.device, dt)
        return y if attention_mask is None else y * attention_mask[..., None].to(w.device, dt)

    @torch.no_grad()
    def generate(self, inputs=None, generation_config=None, logits_processor=None, stopping_criteria=None,
                 synced_gpus=None, streamer=None, **kwargs):
        """generate(input_ids=description ids, prompt_input_ids=transcript ids, ...) -> audio (B, T)
        (modeling_parler_tts.py:2342-2686)."""
        gc, mk = _resolve_generation(self.generation_config, generation_config, kwargs, None)
        seed = mk.pop("seed", None)
        row_offset = mk.pop("row_offset", 0)
        return_codes = mk.pop("return_codes", False)
        input_ids = inputs if inputs is not None else mk.pop("input_ids", None)
        input_values = mk.pop("input_values", None)
        dev = self.device
        attention_mask = mk.pop("attention_mask", None)
        mask_is_trivial = None  # True: every description position is attended (the decoder then runs without the mask)
        enc_out = mk.pop("encoder_outputs", None)
        if enc_out is not None:
            hs = enc_out[0] if isinstance(enc_out, (tuple, list)) else enc_out.last_hidden_state
        else:
            if input_ids is None:
                raise ValueError("generate() needs input_ids (description token ids) or encoder_outputs")
            input_ids = input_ids.to(dev)
            if attention_mask is None:  # _prepare_attention_mask_for_generation: mask only if pad is distinguishable
                pad, eos = gc.pad_token_id, gc.eos_token_id
                if pad is not None and pad != eos and bool((input_ids == pad).any()):
                    attention_mask = (input_ids != pad).long()
                else:
                    attention_mask = torch.ones_like(input_ids)
                    mask_is_trivial = True
            attention_mask = attention_mask.to(dev)
            if mask_is_trivial is None:  # a caller's mask: looked at now, before the encoder is queued
                mask_is_trivial = _drop_if_all_ones(attention_mask) is None
            with _nvtx("ptts.t5_encode"):
                hs = self.encode_text(input_ids, attention_mask)
        B = hs.shape[0]
        nret = _gcv(gc, "num_return_sequences")
        if not _gcv(gc, "do_sample") and nret > 1:
            raise ValueError(f"num_return_sequences has to be 1 when doing greedy search, but is {nret}.")
        cfg = gc.guidance_scale is not None and gc.guidance_scale > 1
        prompt_ids = mk.pop("prompt_input_ids", None)
        prompt_hidden = mk.pop("prompt_hidden_states", None)
        prompt_mask = _drop_if_all_ones(mk.pop("prompt_attention_mask", None))
        dec_ids = mk.pop("decoder_input_ids", None)
        if input_values is not None:
            # audio-prompt continuation (modeling_parler_tts.py:2247-2292): the prompt's codes become decoder_input_ids
            enc_codes = self.audio_encoder.encode(input_values.to(dev)).audio_codes
            frames, bsz, codebooks, seq_len = enc_codes.shape
            if frames != 1:
                raise ValueError(f"Expected 1 frame in the audio code outputs, got {frames} frames. Ensure chunking is "
                                 "disabled by setting `chunk_length=None` in the audio encoder.")
            dec_ids = enc_codes[0].reshape(bsz * self.config.decoder.num_codebooks, seq_len)
        if cfg:  # null conditioning for the second half (modeling_parler_tts.py:2231-2237, 2115-2118)
            hs = torch.cat([hs, torch.zeros_like(hs)], dim=0)
            if attention_mask is not None:
                attention_mask = torch.cat([attention_mask, torch.zeros_like(attention_mask)], dim=0)
            prompt_ids = None if prompt_ids is None else prompt_ids.repeat(2, 1)
            prompt_hidden = None if prompt_hidden is None else prompt_hidden.repeat(2, 1, 1)
            prompt_mask = None if prompt_mask is None else prompt_mask.repeat(2, 1)
        if nret > 1:
            rep = lambda t: None if t is None else t.repeat_interleave(nret, dim=0)  # noqa: E731
            hs, attention_mask, prompt_ids, prompt_hidden, prompt_mask, dec_ids = map(
                rep, (hs, attention_mask, prompt_ids, prompt_hidden, prompt_mask, dec_ids))
        enc = self.project_encoder(hs, attention_mask)
        if attention_mask is not None and not cfg and (mask_is_trivial or _drop_if_all_ones(attention_mask) is None):
            # all ones (the usual equal-length batch, modeling_parler_tts.py's ones_like default): the cross-attention of
            # every decode step would load and test the mask for nothing - 1.259 against 1.203 ms per step at B=64
            # (profiles/r02_e2e_phases_fine_v19.txt)
            attention_mask = None
        prompt_table = None
        if prompt_hidden is not None:
            prompt_hidden, prompt_ids = prompt_hidden.to(dev, torch.float32).contiguous(), None
        elif prompt_ids is not None:
            prompt_ids = prompt_ids.to(dev, torch.int64).contiguous()
            prompt_table = self.prompt_table()
        if streamer is None and not return_codes and self.overlap_decode_chunk_frames > 0 and dec_ids is None:
            streamer = SilentChunks(self.overlap_decode_chunk_frames, self.overlap_decode_ctas)
        codes, seq, st = _generate_codes(self.decoder, enc, attention_mask, prompt_table, prompt_ids, prompt_hidden,
                                         prompt_mask, dec_ids, gc, kwargs, logits_processor, stopping_criteria,
                                         streamer, self.decoder.impl, self.decoder.use_cuda_graph,
                                         _draw_seed() if seed is None else seed, row_offset,
                                         audio_engine=None if return_codes or not isinstance(streamer, AudioChunkStreamer)
                                         else self.audio_encoder.engine(self.decoder.impl))
        if return_codes:
            return codes
        dac = self.audio_encoder.engine(self.decoder.impl)
        special_ids = (gc.bos_token_id if gc.bos_token_id is not None else self.config.decoder.bos_token_id,
                       gc.pad_token_id if gc.pad_token_id is not None else self.config.decoder.pad_token_id,
                       _first_id(gc.eos_token_id) if gc.eos_token_id is not None else self.config.decoder.eos_token_id)
        with _nvtx("ptts.dac_decode"):
            if st.audio_pipeline is not None:  # chunks were decoded during the loop: the remainder, then the whole
                audio = st.audio_pipeline.finish(codes, lambda c: dac.decode_generated(c, *special_ids))
                st.audio_pipeline = None
            else:
                audio = dac.decode_generated(codes, *special_ids)
                if isinstance(streamer, AudioChunkStreamer):  # host-processor loop / audio prompt: one chunk at the end
                    streamer.put_audio(0, audio.cpu() if streamer.to_host else audio)
                    streamer.end()
        if gc.return_dict_in_generate:
            return SimpleNamespace(sequences=audio, audio_codes=codes)
        return audio
