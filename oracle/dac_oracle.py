"""ORACLE (test infrastructure, CPU, torch fp32/fp64) - Descript Audio Codec decode: codes -> waveform.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this.
It is the checker for the CUDA path, never the product path.

The arithmetic lives in a third-party dependency that is NOT under /root/reference: `descript-audio-codec`
(unpinned, /root/reference/setup.py:23).  Call sites that anchor it:
  /root/reference/parler_tts/dac_wrapper/modeling_dac.py:19-23   DAC(n_codebooks, latent_dim, codebook_size)
  /root/reference/parler_tts/dac_wrapper/modeling_dac.py:130     quantizer.from_codes(codes)[0]
  /root/reference/parler_tts/dac_wrapper/modeling_dac.py:131     model.decode(z)
  /root/reference/parler_tts/modeling_parler_tts.py:2654-2680    batched / per-sample ragged dispatch
Published algorithm (Kumar et al. 2023, "High-Fidelity Audio Compression with Improved RVQGAN"; 44.1 kHz config):
  z = sum_i out_proj_i(codebook_i[codes_i])                       RVQ from_codes, codebook dim 8 -> latent 1024
  y = conv7(z) 1024->1536;  4 x { snake; convT(k=2s, stride s, pad ceil(s/2)) C->C/2;
       3 x residual unit(dilation 1,3,9: snake, conv7 dilated, snake, conv1, + skip) }   s = 8, 8, 4, 2
  wav = tanh(conv7(snake(y)))  96 -> 1
  snake(x) = x + sin^2(alpha x) / (alpha + 1e-9); every conv is weight-normalised (w = g * v / ||v||, norm over all
  dims but 0).
State-dict keys are descript's, as stored verbatim under `audio_encoder.model.*` in Parler-TTS checkpoints
(/root/reference/helpers/push_to_hub_scripts/push_dac_to_hub.py:9-13): `quantizer.quantizers.{i}.codebook.weight`,
`quantizer.quantizers.{i}.out_proj.{weight_g,weight_v,bias}`, `decoder.model.0.*` (first conv),
`decoder.model.{1..4}.block.0.alpha`, `.block.1.*` (convT), `.block.{2,3,4}.block.{0,2}.alpha`,
`.block.{2,3,4}.block.{1,3}.*`, `decoder.model.5.alpha`, `decoder.model.6.*`.

Pinned against the arithmetic-equivalent implementation that ships in the installed `transformers`
(models/dac/modeling_dac.py: Snake1d :85-99, ResidualUnit :173-207, DecoderBlock :234-263, from_codes :345-369,
Decoder :405-439) on identical weights: tests/golden/dac_tiny.npz (oracle/gen_golden.py).  No reference-repo golden
vectors exist for it.
"""
import math

import torch
import torch.nn.functional as F

DAC_44K = dict(n_codebooks=9, codebook_size=1024, codebook_dim=8, latent_dim=1024, decoder_dim=1536, rates=(8, 8, 4, 2))


def wn(sd, prefix, dtype=None):
    """Fold weight norm: w = g * v / ||v||, norm over all dims except 0."""
    if prefix + ".weight" in sd:
        w = sd[prefix + ".weight"]
    else:
        g, v = sd[prefix + ".weight_g"], sd[prefix + ".weight_v"]
        nrm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
        w = g * v / nrm
    b = sd.get(prefix + ".bias")
    if dtype is not None:
        w = w.to(dtype)
        b = None if b is None else b.to(dtype)
    return w, b


def snake(x, alpha):
    return x + (alpha + 1e-9).reciprocal() * torch.sin(alpha * x).pow(2)


def random_dac_state(cfg=DAC_44K, seed=0, dtype=torch.float32, alpha_jitter=0.2):
    """Random descript-layout decoder + quantizer weights (weight_g / weight_v form)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def conv(prefix, cout, cin, k, transposed=False, gain=1.0):
        shape = (cin, cout, k) if transposed else (cout, cin, k)
        fan_in = (cout if transposed else cin) * k
        v = torch.randn(*shape, generator=g) / math.sqrt(fan_in)
        sd[prefix + ".weight_v"] = v.to(dtype)
        # g near gain*||v||: activations stay O(1) through ~30 layers (like a trained codec), so that the comparison
        # is well conditioned (snake's sin(alpha x) amplifies rounding noise when |x| grows)
        n = gain * v.reshape(shape[0], -1).norm(dim=1).reshape(-1, 1, 1)
        sd[prefix + ".weight_g"] = (n * (1.0 + 0.1 * torch.randn(shape[0], 1, 1, generator=g))).to(dtype)
        sd[prefix + ".bias"] = (0.02 * torch.randn(cout, generator=g)).to(dtype)

    def alpha(prefix, c):
        sd[prefix + ".alpha"] = (1.0 + alpha_jitter * torch.randn(1, c, 1, generator=g)).abs().add(0.05).to(dtype)

    for i in range(cfg["n_codebooks"]):
        q = f"quantizer.quantizers.{i}"
        sd[q + ".codebook.weight"] = torch.randn(cfg["codebook_size"], cfg["codebook_dim"], generator=g).to(dtype)
        conv(q + ".out_proj", cfg["latent_dim"], cfg["codebook_dim"], 1, gain=1.0 / 3.0)
    ch = cfg["decoder_dim"]
    conv("decoder.model.0", ch, cfg["latent_dim"], 7)
    for bi, s in enumerate(cfg["rates"]):
        p = f"decoder.model.{bi + 1}.block"
        cin, cout = ch // 2 ** bi, ch // 2 ** (bi + 1)
        alpha(p + ".0", cin)
        conv(p + ".1", cout, cin, 2 * s, transposed=True, gain=1.0)
        for ri in range(3):
            r = f"{p}.{ri + 2}.block"
            alpha(r + ".0", cout)
            conv(r + ".1", cout, cout, 7)
            alpha(r + ".2", cout)
            conv(r + ".3", cout, cout, 1, gain=0.3)
    n = len(cfg["rates"])
    cl = ch // 2 ** n
    alpha(f"decoder.model.{n + 1}", cl)
    conv(f"decoder.model.{n + 2}", 1, cl, 7, gain=1.0)
    return sd


def from_codes(sd, codes, cfg=DAC_44K, dtype=torch.float32):
    """codes (B, K, F) int64 -> z (B, latent, F)."""
    z = 0.0
    for i in range(codes.shape[1]):
        q = f"quantizer.quantizers.{i}"
        e = F.embedding(codes[:, i, :], sd[q + ".codebook.weight"].to(dtype)).transpose(1, 2)
        w, b = wn(sd, q + ".out_proj", dtype)
        z = z + F.conv1d(e, w, b)
    return z


def decode_latents(sd, z, cfg=DAC_44K, dtype=torch.float32, taps=None):
    w, b = wn(sd, "decoder.model.0", dtype)
    x = F.conv1d(z.to(dtype), w, b, padding=3)
    if taps is not None:
        taps.append(x)
    for bi, s in enumerate(cfg["rates"]):
        p = f"decoder.model.{bi + 1}.block"
        x = snake(x, sd[p + ".0.alpha"].to(dtype))
        w, b = wn(sd, p + ".1", dtype)
        # no output_padding: an odd stride s yields s*T - 1 samples, as in transformers' DacDecoderBlock (the anchor for the
        # un-vendored descript-audio-codec, models/dac/modeling_dac.py:243-249); the reference's codec has even strides only
        x = F.conv_transpose1d(x, w, b, stride=s, padding=math.ceil(s / 2))
        if taps is not None:
            taps.append(x)
        for ri, dil in enumerate((1, 3, 9)):
            r = f"{p}.{ri + 2}.block"
            y = snake(x, sd[r + ".0.alpha"].to(dtype))
            w, b = wn(sd, r + ".1", dtype)
            y = F.conv1d(y, w, b, dilation=dil, padding=3 * dil)
            y = snake(y, sd[r + ".2.alpha"].to(dtype))
            w, b = wn(sd, r + ".3", dtype)
            y = F.conv1d(y, w, b)
            x = x + y
            if taps is not None:
                taps.append(x)
    n = len(cfg["rates"])
    x = snake(x, sd[f"decoder.model.{n + 1}.alpha"].to(dtype))
    w, b = wn(sd, f"decoder.model.{n + 2}", dtype)
    x = torch.tanh(F.conv1d(x, w, b, padding=3))
    return x  # (B, 1, hop * F)


def decode_codes(sd, codes, cfg=DAC_44K, dtype=torch.float32):
    """DACModel.decode (dac_wrapper/modeling_dac.py:98-134): codes (B,K,F) -> audio (B, 1, T)."""
    return decode_latents(sd, from_codes(sd, codes, cfg, dtype), cfg, dtype)


def decode_generated(sd, codes, bos, pad, eos, codebook_size, cfg=DAC_44K, dtype=torch.float32):
    """modeling_parler_tts.py:2650-2680: batched when no special id survived de-delay, else per sample with invalid
    frames dropped and zero right-padding.  codes (B,K,F) int64 torch -> (B, T)."""
    special = bool(((codes == bos) | (codes == pad) | (codes == eos)).any())
    if not special:
        return decode_codes(sd, codes, cfg, dtype).squeeze(1)
    outs = []
    for b in range(codes.shape[0]):
        keep = (codes[b] >= codebook_size).sum(dim=0) == 0
        if int(keep.sum()) > 0:
            outs.append(decode_codes(sd, codes[b : b + 1][:, :, keep], cfg, dtype)[0, 0])
        else:
            outs.append(torch.zeros(1, dtype=dtype))
    return torch.nn.utils.rnn.pad_sequence(outs, batch_first=True, padding_value=0.0)


# ------------------------------------------------------------------------------------------------------ encode
# Audio-prompt continuation: /root/reference/parler_tts/modeling_parler_tts.py:2247-2292 calls
# DACModel.encode (dac_wrapper/modeling_dac.py:25-96) = descript `DAC.preprocess` (right-pad to a multiple of the hop)
# + `DAC.encode` (encoder conv stack, then residual vector quantisation).  Arithmetic as mirrored by transformers
# models/dac/modeling_dac.py: DacEncoder :442-473, DacEncoderBlock :210-231, DacVectorQuantize :122-170 (factorised,
# l2-normalised nearest neighbour), DacResidualVectorQuantizer.forward :281-343.  Descript key names:
# `encoder.block.0` (conv k7), `encoder.block.{1..4}.block.{0,1,2}.block.{0 alpha,1 conv7,2 alpha,3 conv1}`,
# `encoder.block.{i}.block.3.alpha`, `encoder.block.{i}.block.4` (strided conv k=2s), `encoder.block.5.alpha`,
# `encoder.block.6` (conv k3), `quantizer.quantizers.{i}.in_proj`.
def random_dac_encoder_state(cfg=DAC_44K, seed=0, dtype=torch.float32, alpha_jitter=0.2, encoder_dim=64):
    """Random descript-layout encoder + in_proj weights.  Separate generator: random_dac_state's stream (and every
    fixture regenerated from it) stays as it is."""
    g = torch.Generator().manual_seed(seed + 1000)
    sd = {}

    def conv(prefix, cout, cin, k, gain=1.0):
        v = torch.randn(cout, cin, k, generator=g) / math.sqrt(cin * k)
        sd[prefix + ".weight_v"] = v.to(dtype)
        n = gain * v.reshape(cout, -1).norm(dim=1).reshape(-1, 1, 1)
        sd[prefix + ".weight_g"] = (n * (1.0 + 0.1 * torch.randn(cout, 1, 1, generator=g))).to(dtype)
        sd[prefix + ".bias"] = (0.02 * torch.randn(cout, generator=g)).to(dtype)

    def alpha(prefix, c):
        sd[prefix + ".alpha"] = (1.0 + alpha_jitter * torch.randn(1, c, 1, generator=g)).abs().add(0.05).to(dtype)

    d = encoder_dim
    conv("encoder.block.0", d, 1, 7, gain=2.0)
    strides = tuple(reversed(cfg["rates"]))
    for bi, s in enumerate(strides):
        p = f"encoder.block.{bi + 1}.block"
        for ri in range(3):
            r = f"{p}.{ri}.block"
            alpha(r + ".0", d)
            conv(r + ".1", d, d, 7)
            alpha(r + ".2", d)
            conv(r + ".3", d, d, 1, gain=0.3)
        alpha(p + ".3", d)
        conv(p + ".4", 2 * d, d, 2 * s)
        d *= 2
    n = len(strides)
    alpha(f"encoder.block.{n + 1}", d)
    conv(f"encoder.block.{n + 2}", cfg["latent_dim"], d, 3)
    for i in range(cfg["n_codebooks"]):
        conv(f"quantizer.quantizers.{i}.in_proj", cfg["codebook_dim"], cfg["latent_dim"], 1, gain=1.0)
    return sd


def hop_length(cfg=DAC_44K):
    h = 1
    for s in cfg["rates"]:
        h *= s
    return h


def preprocess(audio, cfg=DAC_44K):
    """descript DAC.preprocess: zero right-pad (B, 1, T) to a multiple of the hop length."""
    hop = hop_length(cfg)
    T = audio.shape[-1]
    return F.pad(audio, (0, math.ceil(T / hop) * hop - T))


def encode_latents(sd, audio, cfg=DAC_44K, dtype=torch.float32):
    """(B, 1, T) -> z (B, latent, T / hop): the encoder conv stack."""
    w, b = wn(sd, "encoder.block.0", dtype)
    x = F.conv1d(audio.to(dtype), w, b, padding=3)
    strides = tuple(reversed(cfg["rates"]))
    for bi, s in enumerate(strides):
        p = f"encoder.block.{bi + 1}.block"
        for ri, dil in enumerate((1, 3, 9)):
            r = f"{p}.{ri}.block"
            y = snake(x, sd[r + ".0.alpha"].to(dtype))
            w, b = wn(sd, r + ".1", dtype)
            y = F.conv1d(y, w, b, dilation=dil, padding=3 * dil)
            y = snake(y, sd[r + ".2.alpha"].to(dtype))
            w, b = wn(sd, r + ".3", dtype)
            x = x + F.conv1d(y, w, b)
        x = snake(x, sd[p + ".3.alpha"].to(dtype))
        w, b = wn(sd, p + ".4", dtype)
        x = F.conv1d(x, w, b, stride=s, padding=math.ceil(s / 2))
    n = len(strides)
    x = snake(x, sd[f"encoder.block.{n + 1}.alpha"].to(dtype))
    w, b = wn(sd, f"encoder.block.{n + 2}", dtype)
    return F.conv1d(x, w, b, padding=1)


def rvq_encode(sd, z, cfg=DAC_44K, dtype=torch.float32, n_quantizers=None, return_margins=False):
    """z (B, latent, F) -> codes (B, K, F) int64: per codebook in_proj -> nearest l2-normalised codebook row (first
    index on ties) -> residual -= out_proj(codebook[code]).  margins = best minus second-best similarity."""
    residual = z.to(dtype)
    codes, margins = [], []
    K = cfg["n_codebooks"] if n_quantizers is None else n_quantizers
    for i in range(K):
        q = f"quantizer.quantizers.{i}"
        w, b = wn(sd, q + ".in_proj", dtype)
        e = F.conv1d(residual, w, b)  # (B, D, F)
        B_, D, Fr = e.shape
        enc = F.normalize(e.permute(0, 2, 1).reshape(B_ * Fr, D))
        cb = F.normalize(sd[q + ".codebook.weight"].to(dtype))
        dist = -(enc.pow(2).sum(1, keepdim=True) - 2 * enc @ cb.t()) + cb.pow(2).sum(1, keepdim=True).t()
        top2 = dist.topk(2, dim=1).values
        margins.append((top2[:, 0] - top2[:, 1]).reshape(B_, Fr))
        idx = dist.max(1)[1].reshape(B_, Fr)
        codes.append(idx)
        zq = F.embedding(idx, sd[q + ".codebook.weight"].to(dtype)).transpose(1, 2)
        zq = e + (zq - e)  # the straight-through expression the reference evaluates in the forward pass (:149)
        w, b = wn(sd, q + ".out_proj", dtype)
        residual = residual - F.conv1d(zq, w, b)
    codes = torch.stack(codes, dim=1)
    return (codes, torch.stack(margins, dim=1)) if return_margins else codes


def encode_audio(sd, audio, cfg=DAC_44K, dtype=torch.float32):
    """DACModel.encode: (B, 1, T) waveform -> codes (B, K, ceil(T / hop))."""
    return rvq_encode(sd, encode_latents(sd, preprocess(audio, cfg), cfg, dtype), cfg, dtype)
