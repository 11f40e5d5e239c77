"""DAC decoder with up-sampling rates other than the 44.1 kHz model's (8, 8, 4, 2): DACModel reads them from the
checkpoint (modeling.py _dac_rates_of); both kernel families against the CPU oracle's waveform."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import parler_tts_b200 as ptts  # noqa: E402
from oracle import dac_oracle  # noqa: E402
from parler_tts_b200 import cabi  # noqa: E402

dev = torch.device("cuda", 0)
for rates in ((4, 2, 8, 2), (2, 2, 2, 2), (8, 4), (16, 2, 2)):
    cfg = dict(n_codebooks=3, codebook_size=32, codebook_dim=4, latent_dim=64, decoder_dim=64 * 2 ** max(len(rates) - 4, 0),
               rates=rates)
    sd = dac_oracle.random_dac_state(cfg, seed=5)
    codes = torch.randint(0, 32, (2, 3, 9), generator=torch.Generator().manual_seed(1))
    want = dac_oracle.decode_codes(sd, codes, cfg).squeeze(1)
    m = ptts.DACModel(ptts.DACConfig(num_codebooks=3, codebook_size=32, latent_dim=64), state_dict=sd).to(dev)
    assert m.rates == rates
    for impl in (cabi.IMPL_SIMPLE, cabi.IMPL_FAST):
        try:
            got = m.engine(impl).decode(codes.to(dev)).squeeze(1).cpu()
            print(rates, "impl", impl, tuple(got.shape), "max abs err", float((got - want).abs().max()))
        except Exception as e:  # noqa: BLE001
            print(rates, "impl", impl, "ERROR", type(e).__name__, str(e)[:200])

# the tcgen05 conv path (44.1 kHz widths: 1024 -> 1536 -> ... -> 96) with the rates permuted
for rates in ((2, 4, 8, 8), (4, 8, 2, 8)):
    cfg = dict(dac_oracle.DAC_44K, rates=rates)
    sd = dac_oracle.random_dac_state(cfg, seed=2)
    codes = torch.randint(0, 1024, (1, 9, 6), generator=torch.Generator().manual_seed(6))
    want = dac_oracle.decode_codes(sd, codes, cfg).squeeze(1)
    m = ptts.DACModel(ptts.DACConfig(), state_dict=sd).to(dev)
    try:
        eng = m.engine(cabi.IMPL_FAST)
        got = eng.decode(codes.to(dev)).squeeze(1).cpu()
        print("44k widths", rates, "fast family, tcgen05 path:", eng.nwc, tuple(got.shape), "max abs err",
              float((got - want).abs().max()))
    except Exception as e:  # noqa: BLE001
        print("44k widths", rates, "ERROR", type(e).__name__, str(e)[:200])
