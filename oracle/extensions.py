"""CPU restatements of the two extensions that have NO counterpart in the reference's code (TEST INFRASTRUCTURE ONLY;
PARITY UNPINNED: the reference cannot arbitrate, the definitions below are this implementation's, stated in
include/mixfe.h and csrc/k9_extensions.cuh):

* weighted_masking      the proposal's activity- and band-weighted masking (PDF p.2 section 2.4); the un-weighted form
                        the reference implements is src/mixing_utils.py:278-306 (oracle: mixfe_oracle.masking_features)
* envelope_compress     compressor with a decoupled smooth peak detector (Giannoulis, Massberg, Reiss, JAES 2012); the
                        reference's memoryless curve is src/mixing_utils.py:435-447 (oracle: mixfe_oracle.compress_static)
"""
from __future__ import annotations

import math
from typing import Dict

import numpy as np
import torch

from . import mixfe_oracle as O


def a_weighting_band_weights(cfg: O.FrontEndConfig = O.FrontEndConfig()) -> torch.Tensor:
    """IEC 61672 A-weighting as a power ratio (max = 1) at the HTK centre frequency of every mel band."""
    top = 2595.0 * math.log10(1.0 + float(cfg.sample_rate // 2) / 700.0)
    mel = np.linspace(0.0, top, cfg.n_mels + 2)[1:-1]
    f = 700.0 * (10.0 ** (mel / 2595.0) - 1.0)
    ra = (12194.0 ** 2 * f ** 4) / ((f ** 2 + 20.6 ** 2) * np.sqrt((f ** 2 + 107.7 ** 2) * (f ** 2 + 737.9 ** 2)) * (f ** 2 + 12194.0 ** 2))
    w = ra ** 2
    return torch.from_numpy(w / w.max())


def weighted_masking(stems: Dict[str, torch.Tensor], cfg: O.FrontEndConfig = O.FrontEndConfig(), beta: float = 0.0,
                     tau: float = 1.0) -> torch.Tensor:
    """dict of 4 x (2,T) -> (4,) fp64:  M_i = sum a_i(t) A(k) sigmoid((beta - Delta_i)/tau) / sum a_i(t) A(k)."""
    fe = O.MelFrontEnd(cfg, torch.float64)
    S = torch.stack([fe.mel_power(stems[n].double()).mean(dim=0) for n in O.STEMS])          # (4, n_mels, F)
    A = a_weighting_band_weights(cfg)[:, None]
    e0 = 1e-6 * (3.0 * cfg.n_fft / 8.0) * cfg.n_freq            # mel energy of a -60 dBFS white-noise frame
    out = []
    for i in range(4):
        others = torch.stack([S[j] for j in range(4) if j != i]).max(dim=0).values
        mu = torch.sigmoid((beta - (S[i] - others)) / tau)
        E = S[i].sum(dim=0)
        a = torch.where(E > 1e-6 * e0, E / (E + e0), torch.zeros_like(E))[None, :]   # under -120 dBFS: not active at all
        den = (a * A).sum()
        out.append((a * A * mu).sum() / den if float(den) > 0 else torch.tensor(0.0, dtype=torch.float64))
    return torch.stack(out)


def envelope_compress(x: torch.Tensor, threshold: float = -20.0, ratio: float = 4.0, attack_ms: float = 5.0,
                      release_ms: float = 50.0, sample_rate: int = 44100) -> torch.Tensor:
    """(C,T) -> (C,T) fp64: p[n] = max(|x[n]|, a_R p[n-1]); s[n] = a_A s[n-1] + (1 - a_A) p[n]; the reference's static curve
    on the level of s; y = x * gain(s)."""
    a_att = math.exp(-1.0 / (attack_ms * 1e-3 * sample_rate)) if attack_ms > 0 else 0.0
    a_rel = math.exp(-1.0 / (release_ms * 1e-3 * sample_rate)) if release_ms > 0 else 0.0
    xs = x.double().numpy()
    out = np.empty_like(xs)
    for c in range(xs.shape[0]):
        p = s = 0.0
        row, dst = xs[c], out[c]
        for n in range(row.shape[0]):
            v = abs(row[n])
            p = v if v > a_rel * p else a_rel * p
            s = a_att * s + (1.0 - a_att) * p
            a = s + 1e-8
            db = 20.0 * math.log10(a)
            if db > threshold:
                db_c = threshold + (db - threshold) / ratio
                dst[n] = row[n] * (10.0 ** (db_c / 20.0)) / a
            else:
                dst[n] = row[n]
    return torch.from_numpy(out)
