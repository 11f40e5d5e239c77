"""DAC (Descript Audio Codec) decode on the CUDA path: codes -> waveform.

Host side of /root/reference/parler_tts/dac_wrapper/modeling_dac.py:98-134 (`DACModel.decode` =
`quantizer.from_codes` + `model.decode`, arithmetic in the un-vendored descript-audio-codec) and of the batched /
ragged dispatch in /root/reference/parler_tts/modeling_parler_tts.py:2650-2680.  Weight-norm is folded once at load
(w = g * v / ||v||) and every codebook is pre-multiplied with its out_proj into a [codebook_size][latent] table; all
per-sample arithmetic runs in libparler_b200.so (ptts_dac_*).
"""
import math

import torch

from . import cabi


def _fold(sd, prefix):
    if prefix + ".weight" in sd:
        w = sd[prefix + ".weight"].float()
    else:
        g, v = sd[prefix + ".weight_g"].float(), sd[prefix + ".weight_v"].float()
        n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
        w = g * v / n
    b = sd.get(prefix + ".bias")
    return w.contiguous(), (None if b is None else b.float().contiguous())


class DacDecoderEngine:
    """Folded fp32 device weights + the launch sequence of the decoder stack."""

    def __init__(self, sd, device, rates=(8, 8, 4, 2), impl=cabi.IMPL_FAST):
        dev = torch.device(device)
        sd = {k: v.detach().to(dev) for k, v in sd.items() if k.startswith(("quantizer.", "decoder."))}
        self.device, self.impl, self.rates = dev, impl, tuple(rates)
        self.hop = int(math.prod(rates))
        n_q = 0
        while f"quantizer.quantizers.{n_q}.codebook.weight" in sd:
            n_q += 1
        tabs = []
        for i in range(n_q):
            q = f"quantizer.quantizers.{i}"
            w, b = _fold(sd, q + ".out_proj")  # [latent][dim][1]
            t = sd[q + ".codebook.weight"].float() @ w[:, :, 0].t()
            tabs.append(t if b is None else t + b)
        self.tables = torch.stack(tabs).contiguous()  # [K][codebook_size][latent]
        self.n_codebooks, self.codebook_size, self.latent = self.tables.shape
        self.conv_in = _fold(sd, "decoder.model.0")
        self.blocks = []
        for bi, s in enumerate(self.rates):
            p = f"decoder.model.{bi + 1}.block"
            blk = dict(stride=s, alpha=sd[p + ".0.alpha"].float().reshape(-1).contiguous(), convt=_fold(sd, p + ".1"),
                       res=[])
            for ri, dil in enumerate((1, 3, 9)):
                r = f"{p}.{ri + 2}.block"
                blk["res"].append(dict(dil=dil, a1=sd[r + ".0.alpha"].float().reshape(-1).contiguous(),
                                       c1=_fold(sd, r + ".1"), a2=sd[r + ".2.alpha"].float().reshape(-1).contiguous(),
                                       c2=_fold(sd, r + ".3")))
            self.blocks.append(blk)
        n = len(self.rates)
        self.alpha_out = sd[f"decoder.model.{n + 1}.alpha"].float().reshape(-1).contiguous()
        self.conv_out = _fold(sd, f"decoder.model.{n + 2}")
        # fast family: channels-last activations + tensor-core implicit GEMM (needs channel counts that tile by 32)
        chans = [self.latent, self.conv_in[0].shape[0]] + [b["convt"][0].shape[1] for b in self.blocks]
        self.nwc = (impl == cabi.IMPL_FAST and all(c % 32 == 0 for c in chans) and self.conv_out[0].shape[0] == 1
                    and self.latent % 4 == 0 and all(s >= 2 for s in self.rates))
        if self.nwc:
            self._prep_nwc()

    # ---- fast family: pre-split bf16 weight planes (w = hi + mid), tap-major, channel-contiguous --------------
    @staticmethod
    def _split(w):
        hi = w.to(torch.bfloat16)
        mid = (w - hi.float()).to(torch.bfloat16)
        return hi.contiguous(), mid.contiguous()

    @staticmethod
    def _tile(hi, mid):
        """[taps][N][Cin] hi / mid planes -> the pre-tiled stage image of the warp-specialised tcgen05 kernel:
        [n_tile][Cin/32][taps][plane][4][bn][8] (one contiguous block per pipeline stage; include/parler_b200.h)."""
        taps, n, cin = hi.shape
        bn = int(cabi.load().ptts_dac_tile_n(n))
        nt = (n + bn - 1) // bn
        w = torch.stack([hi, mid])  # plane, tap, n, ci
        if nt * bn != n:
            w = torch.nn.functional.pad(w, (0, 0, 0, nt * bn - n))
        w = w.view(2, taps, nt, bn, cin // 32, 4, 8).permute(2, 4, 1, 0, 5, 3, 6)  # nt, chunk, tap, plane, kc, n, e
        return w.contiguous(), None

    def _planes(self, w):
        hi, mid = self._split(w)
        return self._tile(hi, mid)

    def _prep_nwc(self):
        def conv(wb):  # [Cout][Cin][k] -> [k][Cout][Cin]
            w, b = wb
            return self._planes(w.permute(2, 0, 1).contiguous()) + (b, w.shape[0], w.shape[2])

        def convt(wb, s):  # [Cin][Cout][2s] -> [2][s*Cout][Cin], index (tap, ph*Cout + co, ci) = w[ci][co][ph + tap*s]
            w, b = wb
            cin, cout, _ = w.shape
            w2 = w.permute(2, 1, 0).reshape(2, s * cout, cin).contiguous()
            return self._planes(w2) + (b, cout)

        self.f_conv_in = conv(self.conv_in)
        self.f_blocks = []
        for blk in self.blocks:
            fb = dict(stride=blk["stride"], alpha=blk["alpha"], convt=convt(blk["convt"], blk["stride"]), res=[])
            for r in blk["res"]:
                fb["res"].append(dict(dil=r["dil"], a1=r["a1"], c1=conv(r["c1"]), a2=r["a2"], c2=conv(r["c2"])))
            self.f_blocks.append(fb)
        w, b = self.conv_out
        self.f_conv_out = (w[0].t().contiguous(), b)  # [7][Cin]

    def dequant_nwc(self, codes):
        """codes (B, K, F) int64 -> z (B, F, latent) fp32, channels-last."""
        codes = codes.to(self.device, torch.int64).contiguous()
        B, K, F = codes.shape
        z = torch.empty(B, F, self.latent, device=self.device, dtype=torch.float32)
        cabi.call("ptts_dac_dequant_nwc", cabi.ptr(codes), B, K, F, K * F, F, cabi.ptr(self.tables),
                  self.codebook_size, self.latent, cabi.ptr(z), cabi.stream_ptr())
        return z

    PAD_LO, PAD_HI = 8, 264  # PTTS_DAC_PLANE_PAD_LO / _HI of include/parler_b200.h

    def _new_planes(self, B, C, T):
        """Uninitialised activation planes [B][2][C/8][PAD_LO + T + PAD_HI][8] bf16 with zeroed pad rows."""
        p = torch.empty(B, 2, C // 8, self.PAD_LO + T + self.PAD_HI, 8, device=self.device, dtype=torch.bfloat16)
        p[:, :, :, : self.PAD_LO].zero_()
        p[:, :, :, self.PAD_LO + T :].zero_()
        return p

    def _conv_px(self, T, wt, dil=1, x=None, xp=None, alpha=None, residual=None, planes_alpha=None, to_planes=False):
        """conv1d over pre-tiled weights: fp32 (x, Snake alpha) or planes (xp) in, fp32 or planes (activated with
        planes_alpha for the consumer) out."""
        w, _, bias, cout, k = wt
        B = (x if x is not None else xp).shape[0]
        cin = x.shape[2] if x is not None else xp.shape[2] * 8
        y = yp = None
        if to_planes:
            yp = self._new_planes(B, cout, T)
        else:
            y = torch.empty(B, T, cout, device=self.device, dtype=torch.float32)
        cabi.call("ptts_dac_conv1d_px", cabi.ptr(x), cabi.ptr(xp), B, T, cin, cabi.ptr(w), cabi.ptr(bias), cout, k, dil,
                  cabi.ptr(alpha), cabi.ptr(residual), cabi.ptr(y), cabi.ptr(yp), cabi.ptr(planes_alpha),
                  cabi.stream_ptr())
        return yp if to_planes else y

    def _convt_px(self, T, wt, stride, xp):
        w, _, bias, cout = wt
        B, cin = xp.shape[0], xp.shape[2] * 8
        y = torch.empty(B, T * stride, cout, device=self.device, dtype=torch.float32)
        cabi.call("ptts_dac_convtr1d_px", None, cabi.ptr(xp), B, T, cin, cabi.ptr(w), cabi.ptr(bias), cout, stride,
                  None, cabi.ptr(y), cabi.stream_ptr())
        return y

    def _decode_latents_planes(self, z):
        """Pre-tiled path: every Snake that feeds a k=1 conv or a transposed conv is applied by the producing conv's
        epilogue, which writes hi/mid bf16 planes instead of fp32 (modeling_dac / dac.model.dac.Decoder order kept)."""
        B, T, _ = z.shape
        blocks = self.f_blocks
        xp = self._conv_px(T, self.f_conv_in, x=z, planes_alpha=blocks[0]["alpha"], to_planes=True)
        x = None
        for i, blk in enumerate(blocks):
            x = self._convt_px(T, blk["convt"], blk["stride"], xp)
            T *= blk["stride"]
            xp = None
            for j, r in enumerate(blk["res"]):
                yp = self._conv_px(T, r["c1"], dil=r["dil"], x=x, alpha=r["a1"], planes_alpha=r["a2"], to_planes=True)
                if j == len(blk["res"]) - 1 and i + 1 < len(blocks):  # feeds the next block's transposed conv
                    xp = self._conv_px(T, r["c2"], xp=yp, residual=x, planes_alpha=blocks[i + 1]["alpha"], to_planes=True)
                    x = None
                else:
                    x = self._conv_px(T, r["c2"], xp=yp, residual=x)
                del yp
        return x

    def decode_latents_nwc(self, z):
        x = self._decode_latents_planes(z)
        B, T, Cin = x.shape
        w, b = self.f_conv_out
        out = torch.empty(B, 1, T, device=x.device, dtype=torch.float32)
        cabi.call("ptts_dac_conv_out_nwc", cabi.ptr(x), B, T, Cin, cabi.ptr(w), cabi.ptr(b), cabi.ptr(self.alpha_out),
                  cabi.ptr(out), cabi.stream_ptr())
        return out

    # ---- single launches -------------------------------------------------------------------------------------
    def _conv(self, x, wb, dil=1, alpha=None, residual=None, tanh=False):
        w, b = wb
        B, Cin, T = x.shape
        Cout, _, k = w.shape
        y = torch.empty(B, Cout, T, device=x.device, dtype=torch.float32)
        cabi.call("ptts_dac_conv1d", cabi.ptr(x), B, Cin, T, cabi.ptr(w), cabi.ptr(b), Cout, k, dil, cabi.ptr(alpha),
                  cabi.ptr(residual), int(tanh), cabi.ptr(y), self.impl, cabi.stream_ptr())
        return y

    def _convt(self, x, wb, stride, alpha):
        w, b = wb  # [Cin][Cout][2*stride]
        B, Cin, T = x.shape
        Cout = w.shape[1]
        y = torch.empty(B, Cout, T * stride, device=x.device, dtype=torch.float32)
        cabi.call("ptts_dac_convtr1d", cabi.ptr(x), B, Cin, T, cabi.ptr(w), cabi.ptr(b), Cout, stride, cabi.ptr(alpha),
                  cabi.ptr(y), self.impl, cabi.stream_ptr())
        return y

    def dequant(self, codes):
        """codes (B, K, F) int64 -> z (B, latent, F) fp32  (quantizer.from_codes)."""
        codes = codes.to(self.device, torch.int64).contiguous()
        B, K, F = codes.shape
        z = torch.empty(B, self.latent, F, device=self.device, dtype=torch.float32)
        cabi.call("ptts_dac_dequant", cabi.ptr(codes), B, K, F, K * F, F, cabi.ptr(self.tables), None,
                  self.codebook_size, self.latent, cabi.ptr(z), cabi.stream_ptr())
        return z

    def decode_latents(self, z):
        x = self._conv(z, self.conv_in)
        for blk in self.blocks:
            x = self._convt(x, blk["convt"], blk["stride"], blk["alpha"])
            for r in blk["res"]:
                y = self._conv(x, r["c1"], dil=r["dil"], alpha=r["a1"])
                x = self._conv(y, r["c2"], alpha=r["a2"], residual=x)
        return self._conv(x, self.conv_out, alpha=self.alpha_out, tanh=True)  # (B, 1, hop*F)

    def decode(self, codes, max_frames_per_call=32 * 1024):
        """codes (B, K, F) -> audio (B, 1, hop*F); the batch is chunked to bound activation memory."""
        B, K, F = codes.shape
        if F == 0:
            return torch.zeros(B, 1, 0, device=self.device)
        per = max(1, max_frames_per_call // F)
        per = -(-B // -(-B // per))  # equal-sized chunks (64 utterances -> 32 + 32, not 38 + 26)
        if self.nwc:
            outs = [self.decode_latents_nwc(self.dequant_nwc(codes[i : i + per])) for i in range(0, B, per)]
        else:
            outs = [self.decode_latents(self.dequant(codes[i : i + per])) for i in range(0, B, per)]
        return torch.cat(outs, dim=0) if len(outs) > 1 else outs[0]

    # frames of context on either side that an output frame of the 44 kHz decoder can see: first conv (k7) 3 frames +
    # block 1 (stride 8; residual units k7 with dilations 1, 3, 9 = 39 samples, transposed conv 1) 7 frames, the later
    # blocks add less than one frame each at their rates -> 10; 16 leaves a margin and keeps chunks aligned
    STREAM_HALO = 16

    def decode_chunks(self, codes, chunk_frames=128, halo=None):
        """Streaming form of `decode`: yields (first_frame, audio (B, 1, hop*n)) for
        consecutive chunks of `chunk_frames` frames.  Every chunk is decoded with `halo` frames of real context on
        both sides (zero padding only at the true ends, exactly like the one-shot decode), so the concatenation of
        the chunks equals `decode(codes)`; a consumer can start playing / copying chunk i while chunk i+1 decodes."""
        B, K, F = codes.shape
        halo = self.STREAM_HALO if halo is None else int(halo)
        hop = 1
        for r in self.rates:
            hop *= r
        for f0 in range(0, F, chunk_frames):
            f1 = min(F, f0 + chunk_frames)
            lo, hi = max(0, f0 - halo), min(F, f1 + halo)
            wav = self.decode(codes[:, :, lo:hi].contiguous())
            yield f0, wav[:, :, (f0 - lo) * hop : (f1 - lo) * hop]

    def decode_generated(self, codes, bos, pad, eos):
        """modeling_parler_tts.py:2650-2680: batched decode when no special id survived the de-delay, else per sample
        with invalid frames dropped (ptts_filter_frames) and zero right-padding.  Returns (B, T) fp32."""
        codes = codes.to(self.device, torch.int64).contiguous()
        B, K, F = codes.shape
        special = bool(((codes == bos) | (codes == pad) | (codes == eos)).any().item())
        if not special:
            if int(codes.max().item()) >= self.codebook_size or int(codes.min().item()) < 0:
                # the reference's codebook nn.Embedding raises here; fail just as loudly instead of decoding garbage
                raise IndexError("audio code outside [0, codebook_size): index out of range in the DAC codebook")
            return self.decode(codes).squeeze(1)
        packed = torch.empty_like(codes)
        counts = torch.zeros(B, device=self.device, dtype=torch.int32)
        cabi.call("ptts_filter_frames", cabi.ptr(codes), B, K, F, self.codebook_size, cabi.ptr(packed),
                  cabi.ptr(counts), cabi.stream_ptr())
        counts = counts.cpu().tolist()
        outs = []
        for b in range(B):
            if counts[b] > 0:
                outs.append(self.decode(packed[b : b + 1, :, : counts[b]].contiguous())[0, 0])
            else:
                outs.append(torch.zeros(1, device=self.device))
        return torch.nn.utils.rnn.pad_sequence(outs, batch_first=True, padding_value=0.0)


class DacEncoderEngine:
    """Waveform -> codes on the CUDA path (audio-prompt continuation): host side of
    /root/reference/parler_tts/dac_wrapper/modeling_dac.py:25-96 (`DACModel.encode` = descript `DAC.preprocess` +
    `DAC.encode`).  Descript key names: `encoder.block.0` (conv k7), `encoder.block.{1..n}.block.{0,1,2}` (residual
    units), `.block.3.alpha`, `.block.4` (strided conv), `encoder.block.{n+1}.alpha`, `encoder.block.{n+2}` (conv k3),
    `quantizer.quantizers.{i}.{in_proj,out_proj,codebook}`.  Weight norm is folded once at load."""

    def __init__(self, sd, device, rates=(8, 8, 4, 2), impl=cabi.IMPL_FAST):
        dev = torch.device(device)
        sd = {k: v.detach().to(dev) for k, v in sd.items() if k.startswith(("quantizer.", "encoder."))}
        if "encoder.block.0.weight_v" not in sd and "encoder.block.0.weight" not in sd:
            raise cabi.ParlerB200Error("this DAC checkpoint holds no encoder weights (encoder.block.*): audio prompts "
                                       "cannot be encoded")
        self.device, self.impl = dev, impl
        self.strides = tuple(reversed(tuple(rates)))  # the encoder mirrors the decoder's up-sampling rates
        self.hop = int(math.prod(rates))
        alpha = lambda k: sd[k].float().reshape(-1).contiguous()  # noqa: E731
        self.conv_in = _fold(sd, "encoder.block.0")
        self.blocks = []
        for bi, s in enumerate(self.strides):
            p = f"encoder.block.{bi + 1}.block"
            blk = dict(stride=s, res=[], alpha=alpha(p + ".3.alpha"), down=_fold(sd, p + ".4"))
            for ri, dil in enumerate((1, 3, 9)):
                r = f"{p}.{ri}.block"
                blk["res"].append(dict(dil=dil, a1=alpha(r + ".0.alpha"), c1=_fold(sd, r + ".1"),
                                       a2=alpha(r + ".2.alpha"), c2=_fold(sd, r + ".3")))
            self.blocks.append(blk)
        n = len(self.strides)
        self.alpha_out = alpha(f"encoder.block.{n + 1}.alpha")
        self.conv_out = _fold(sd, f"encoder.block.{n + 2}")
        K = 0
        while f"quantizer.quantizers.{K}.codebook.weight" in sd:
            K += 1
        w_in, b_in, w_out, b_out, cb = [], [], [], [], []
        for i in range(K):
            q = f"quantizer.quantizers.{i}"
            w, b = _fold(sd, q + ".in_proj")   # [D][C][1]
            w_in.append(w[:, :, 0])
            b_in.append(b if b is not None else torch.zeros(w.shape[0], device=dev))
            w, b = _fold(sd, q + ".out_proj")  # [C][D][1]
            w_out.append(w[:, :, 0])
            b_out.append(b if b is not None else torch.zeros(w.shape[0], device=dev))
            cb.append(sd[q + ".codebook.weight"].float())
        self.n_codebooks = K
        self.w_in, self.b_in = torch.stack(w_in).contiguous(), torch.stack(b_in).contiguous()
        self.w_out, self.b_out = torch.stack(w_out).contiguous(), torch.stack(b_out).contiguous()
        self.codebook = torch.stack(cb).contiguous()
        self.codebook_n = torch.nn.functional.normalize(self.codebook, dim=-1).contiguous()  # F.normalize(codebook)
        self.codebook_size, self.codebook_dim = self.codebook.shape[1:]
        self.latent = self.w_in.shape[2]

    def _strided(self, x, wb, stride, pad, alpha=None):
        w, b = wb
        B, Cin, T = x.shape
        Cout, _, k = w.shape
        Tout = (T + 2 * pad - k) // stride + 1
        y = torch.empty(B, Cout, Tout, device=x.device, dtype=torch.float32)
        cabi.call("ptts_dac_conv1d_strided", cabi.ptr(x), B, Cin, T, cabi.ptr(w), cabi.ptr(b), Cout, k, stride, pad,
                  cabi.ptr(alpha), cabi.ptr(y), cabi.stream_ptr())
        return y

    def _conv(self, x, wb, dil=1, alpha=None, residual=None):
        w, b = wb
        B, Cin, T = x.shape
        Cout, _, k = w.shape
        y = torch.empty(B, Cout, T, device=x.device, dtype=torch.float32)
        cabi.call("ptts_dac_conv1d", cabi.ptr(x), B, Cin, T, cabi.ptr(w), cabi.ptr(b), Cout, k, dil, cabi.ptr(alpha),
                  cabi.ptr(residual), 0, cabi.ptr(y), self.impl, cabi.stream_ptr())
        return y

    def encode_latents(self, audio):
        """audio (B, 1, T) fp32, T a multiple of the hop -> z (B, latent, T / hop)."""
        x = self._strided(audio, self.conv_in, 1, 3)
        for blk in self.blocks:
            for r in blk["res"]:
                y = self._conv(x, r["c1"], dil=r["dil"], alpha=r["a1"])
                x = self._conv(y, r["c2"], alpha=r["a2"], residual=x)
            s = blk["stride"]
            x = self._strided(x, blk["down"], s, math.ceil(s / 2), alpha=blk["alpha"])
        return self._strided(x, self.conv_out, 1, 1, alpha=self.alpha_out)

    def quantize(self, z, n_quantizers=None):
        """z (B, latent, F) -> codes (B, K, F) int64."""
        z = z.contiguous()
        B, C, F = z.shape
        K = self.n_codebooks if n_quantizers is None else min(int(n_quantizers), self.n_codebooks)
        codes = torch.empty(B, K, F, device=z.device, dtype=torch.int64)
        cabi.call("ptts_dac_vq_encode", cabi.ptr(z), B, C, F, K, self.codebook_dim, self.codebook_size,
                  cabi.ptr(self.w_in), cabi.ptr(self.b_in), cabi.ptr(self.codebook), cabi.ptr(self.codebook_n),
                  cabi.ptr(self.w_out), cabi.ptr(self.b_out), cabi.ptr(codes), cabi.stream_ptr())
        return codes

    def encode(self, audio, n_quantizers=None):
        """audio (B, 1, T) -> codes (B, K, ceil(T / hop)); zero right-padding to a hop multiple like DAC.preprocess."""
        if self.device.type != "cuda":
            raise cabi.ParlerB200Error("the DAC encoder runs on the CUDA path only (no CPU fallback)")
        a = audio.to(self.device, torch.float32)
        T = a.shape[-1]
        a = torch.nn.functional.pad(a, (0, -(-T // self.hop) * self.hop - T)).contiguous()
        return self.quantize(self.encode_latents(a), n_quantizers)
