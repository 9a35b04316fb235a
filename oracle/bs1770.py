"""TEST INFRASTRUCTURE ONLY -- CPU oracle of the K-weighted, gated loudness extension.

PARITY UNPINNED: the reference computes loudness as the un-weighted, un-gated
-0.691 + 10 log10(mean x^2 + 1e-10) (src/mixing_utils.py:311-318); K-weighting and gating appear only in the
proposal (Mixing_Style_Representation_Learning_Proposal.pdf p.2) and, off the hot path, through the un-vendored
`pyloudnorm` package (src/style_transfer_eval.py:246-248; not in requirements.txt, no pinned version).  This file
restates the published algorithm, ITU-R BS.1770-4:
  * K-weighting = high-shelf "pre-filter" followed by the RLB high-pass.  Coefficients for an arbitrary sample rate
    come from the analog prototype behind the Recommendation's 48 kHz table (bilinear transform with pre-warping,
    the derivation used by libebur128); `k_weighting_sos(48000)` reproduces the table of BS.1770-4 Annex 1 to 1e-9,
    which tests/test_oracle_golden.py checks -- that table is the only golden vector this extension has.
  * gating blocks of 400 ms overlapping by 75 %, absolute gate at -70 LUFS, relative gate 10 LU below the loudness
    of the blocks that pass the absolute gate; channel weights 1.0 for left and right.
  Block bounds and the >= / > comparisons follow pyloudnorm's meter (Meter.integrated_loudness), the implementation
  the reference's evaluation script would call.
"""
from __future__ import annotations

import math

import numpy as np

ITU_48K_SHELF = ([1.53512485958697, -2.69169618940638, 1.19839281085285], [1.0, -1.69065929318241, 0.73248077421585])
ITU_48K_HIGHPASS = ([1.0, -2.0, 1.0], [1.0, -1.99004745483398, 0.99007225036621])


def k_weighting_sos(fs: float) -> np.ndarray:
    """(2, 6) second-order sections b0 b1 b2 1 a1 a2: high shelf, then high-pass."""
    f0, G, Q = 1681.974450955533, 3.999843853973347, 0.7071752369554196
    K = math.tan(math.pi * f0 / fs)
    Vh = 10.0 ** (G / 20.0)
    Vb = Vh ** 0.4996667741545416
    a0 = 1.0 + K / Q + K * K
    shelf = [(Vh + Vb * K / Q + K * K) / a0, 2.0 * (K * K - Vh) / a0, (Vh - Vb * K / Q + K * K) / a0,
             1.0, 2.0 * (K * K - 1.0) / a0, (1.0 - K / Q + K * K) / a0]
    f0, Q = 38.13547087602444, 0.5003270373238773
    K = math.tan(math.pi * f0 / fs)
    a0 = 1.0 + K / Q + K * K
    hp = [1.0, -2.0, 1.0, 1.0, 2.0 * (K * K - 1.0) / a0, (1.0 - K / Q + K * K) / a0]
    return np.array([shelf, hp], dtype=np.float64)


def _sosfilt(sos: np.ndarray, x: np.ndarray) -> np.ndarray:
    from scipy.signal import sosfilt
    return sosfilt(sos, x.astype(np.float64), axis=-1)


def sub_block_len(fs: int) -> int:
    if fs % 10:
        raise ValueError("sample rate must be a multiple of 10 Hz (100 ms steps of whole samples)")
    return fs // 10


def integrated_loudness(x: np.ndarray, fs: int) -> float:
    """x: (channels, T) -> integrated loudness in LUFS (-inf when no block passes the gates)."""
    x = np.atleast_2d(np.asarray(x))
    sub = sub_block_len(fs)
    T = x.shape[-1]
    if T < 4 * sub:
        raise ValueError("Audio must have length greater than the block size")
    y = _sosfilt(k_weighting_sos(fs), x)
    n_blocks = (T - 4 * sub) // sub + 1
    z = np.stack([[np.mean(y[c, j * sub:j * sub + 4 * sub] ** 2) for j in range(n_blocks)] for c in range(x.shape[0])])
    with np.errstate(divide="ignore"):
        l = -0.691 + 10.0 * np.log10(z.sum(axis=0))
        keep = l >= -70.0
        if not keep.any():
            return float("-inf")
        gamma_r = -0.691 + 10.0 * np.log10(z[:, keep].mean(axis=1).sum()) - 10.0
        keep = (l > gamma_r) & (l > -70.0)
        if not keep.any():
            return float("-inf")
        return float(-0.691 + 10.0 * np.log10(z[:, keep].mean(axis=1).sum()))


def stems_gated_loudness(stems, fs: int = 44100) -> np.ndarray:
    """dict of 4 x (2,T) torch/numpy stems -> [vocals, bass, drums, other, mixture] LUFS; the mixture is the fp32
    sum in dict order, as src/mixing_utils.py:85 forms it."""
    order = ("vocals", "bass", "drums", "other")
    arrs = [np.asarray(stems[k], dtype=np.float32) for k in order]
    mix = np.zeros_like(arrs[0])
    for a in arrs:
        mix = (mix + a).astype(np.float32)
    return np.array([integrated_loudness(a, fs) for a in arrs + [mix]], dtype=np.float64)


def short_term_loudness(x: np.ndarray, fs: int) -> np.ndarray:
    """x: (channels, T) -> short-term loudness in LUFS of every 3 s window, hop 100 ms (EBU Tech 3341)."""
    x = np.atleast_2d(np.asarray(x))
    sub = sub_block_len(fs)
    T = x.shape[-1]
    if T < 30 * sub:
        raise ValueError("short-term loudness needs at least 3 s of audio")
    y = _sosfilt(k_weighting_sos(fs), x)
    n_win = (T - 30 * sub) // sub + 1
    z = np.stack([[np.mean(y[c, j * sub:j * sub + 30 * sub] ** 2) for j in range(n_win)] for c in range(x.shape[0])])
    with np.errstate(divide="ignore"):
        return -0.691 + 10.0 * np.log10(z.sum(axis=0))


def short_term_iqr(x: np.ndarray, fs: int) -> float:
    """The proposal's DR_IQR = IQR(L_short(t)) (PDF p.2 section 2.1; PARITY UNPINNED like the rest of this file: the
    proposal fixes neither window, gate nor quantile rule).  Windows under -70 LUFS are dropped; q75 - q25 with
    numpy's default linear interpolation; 0 when fewer than two windows remain."""
    l = short_term_loudness(x, fs)
    l = l[l >= -70.0]
    if l.size < 2:
        return 0.0
    q25, q75 = np.quantile(l, [0.25, 0.75])
    return float(q75 - q25)


def stems_short_term_iqr(stems, fs: int = 44100) -> np.ndarray:
    """dict of 4 x (2,T) stems -> [vocals, bass, drums, other, mixture] IQR in LU (mixture as in stems_gated_loudness)."""
    order = ("vocals", "bass", "drums", "other")
    arrs = [np.asarray(stems[k], dtype=np.float32) for k in order]
    mix = np.zeros_like(arrs[0])
    for a in arrs:
        mix = (mix + a).astype(np.float32)
    return np.array([short_term_iqr(a, fs) for a in arrs + [mix]], dtype=np.float64)
