"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's MultiResolutionSTFTLoss (src/loss.py:332-448), the
checker of kernel K8.  Pinned against the unmodified reference class by tests/golden/make_golden_mrstft.py (values
and gradients in tests/golden/mrstft_outputs.npz; tests/test_oracle_golden_mrstft.py)."""
from __future__ import annotations

from typing import Sequence

import torch


def stft_mag(x: torch.Tensor, n_fft: int, hop: int, win: int) -> torch.Tensor:
    """src/loss.py:354-383 followed by torch.abs (:436-437): (B,C,T) or (C,T) -> (B*C, n_fft/2+1, frames)."""
    if x.ndim == 2:
        x = x.unsqueeze(0)
    B, C, T = x.shape
    w = torch.hann_window(win, dtype=x.dtype, device=x.device)
    spec = torch.stft(x.reshape(B * C, T), n_fft=n_fft, hop_length=hop, win_length=win, window=w, return_complex=True)
    return spec.abs()


def resolution_terms(x: torch.Tensor, y: torch.Tensor, n_fft: int, hop: int, win: int):
    """(spectral convergence :395, log-magnitude L1 :408) of one resolution."""
    xm, ym = stft_mag(x, n_fft, hop, win), stft_mag(y, n_fft, hop, win)
    sc = torch.norm(ym - xm, p="fro") / (torch.norm(ym, p="fro") + 1e-8)
    lm = torch.nn.functional.l1_loss(torch.log(xm + 1e-5), torch.log(ym + 1e-5))
    return sc, lm


def mrstft_loss(x: torch.Tensor, y: torch.Tensor, fft_sizes: Sequence[int] = (1024, 2048, 512),
                hop_sizes: Sequence[int] = (256, 512, 128), win_sizes: Sequence[int] = (1024, 2048, 512)) -> torch.Tensor:
    """src/loss.py:410-448: mean over resolutions of sc + lm."""
    total = 0.0
    for n_fft, hop, win in zip(fft_sizes, hop_sizes, win_sizes):
        sc, lm = resolution_terms(x, y, n_fft, hop, win)
        total = total + sc + lm
    return total / len(fft_sizes)


def pair(seed: int, B: int, C: int, T: int):
    """Deterministic (prediction, target) pair: band-limited noise + tones, the prediction a filtered, rescaled copy."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(T, dtype=torch.float32) / 44100.0
    y = 0.05 * torch.randn(B, C, T, generator=g)
    for c in range(C):
        y[:, c] += 0.2 * torch.sin(2 * torch.pi * (110.0 * (c + 1)) * t) * (0.5 + 0.5 * torch.rand(B, 1, generator=g))
    x = 0.8 * y + 0.1 * torch.roll(y, 3, dims=-1) + 0.02 * torch.randn(B, C, T, generator=g)
    return x, y
