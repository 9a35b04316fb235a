"""GPU: K7 (K-weighted, gated loudness; extension, parity unpinned by the reference) against the BS.1770-4 oracle.
Tolerance 1e-3 LU (fp32 output of an fp64 pipeline; the oracle's scipy sosfilt is the same DF2T arithmetic)."""
import numpy as np
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import bs1770 as B
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("T", [17640, 44100, 100001, 441000])
def test_gated_loudness_vs_oracle(cuda_device, T):
    clip = O.musical_clip(T % 97, T)
    clip["bass"] = clip["bass"] * 0.05                       # spread the levels
    half = T // 2
    clip["drums"] = torch.cat([clip["drums"][:, :half], clip["drums"][:, half:] * 1e-3], dim=1)   # relative gate bites
    fe = pkg.MixingFeatureExtractor()
    n0 = pkg.launch_count()
    got = fe.gated_loudness({n: v.to(cuda_device) for n, v in clip.items()})
    assert pkg.launch_count() == n0 + 2
    assert got.shape == (5,) and got.dtype == torch.float32
    ref = B.stems_gated_loudness({n: v.numpy() for n, v in clip.items()}, 44100)
    assert np.abs(got.cpu().numpy().astype(np.float64) - ref).max() < 1e-3, (got.cpu().numpy(), ref)


def test_batched_equals_per_clip_and_silence_is_minus_inf(cuda_device):
    fe = pkg.MixingFeatureExtractor()
    T = 60000
    clips = [O.musical_clip(30 + i, T) for i in range(3)]
    clips[1]["other"] = torch.zeros(2, T)
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    got = fe.gated_loudness(batch)
    assert got.shape == (3, 5)
    assert got[1, 3] == float("-inf")
    for i, c in enumerate(clips):
        one = fe.gated_loudness({n: v.to(cuda_device) for n, v in c.items()})
        assert torch.equal(one, got[i])
    again = fe.gated_loudness(batch)
    assert torch.equal(again, got)                            # fixed-order reductions


def test_too_short_raises(cuda_device):
    fe = pkg.MixingFeatureExtractor()
    with pytest.raises(ValueError, match="block size"):
        fe.gated_loudness({n: torch.zeros(2, 17639, device=cuda_device) for n in O.STEMS})


def test_other_sample_rate(cuda_device):
    fs, T = 48000, 96000
    fe = pkg.MixingFeatureExtractor(sample_rate=fs)
    clip = O.musical_clip(3, T, sample_rate=fs)
    got = fe.gated_loudness({n: v.to(cuda_device) for n, v in clip.items()})
    ref = B.stems_gated_loudness({n: v.numpy() for n, v in clip.items()}, fs)
    assert np.abs(got.cpu().numpy().astype(np.float64) - ref).max() < 1e-3


@pytest.mark.parametrize("seed,T,off", [(1, 17640, 1), (2, 17641, 3), (3, 22049, 2), (4, 35279, 1), (5, 35281, 0),
                                        (6, 61237, 3), (7, 130003, 2)])
def test_awkward_lengths_and_unaligned_views(cuda_device, seed, T, off):
    """Lengths around the 100 ms sub-block edges (4410 samples) and the 8-sub-block chunk edge (35280), stems passed as
    views that start `off` floats into a wider tensor."""
    clip = O.musical_clip(500 + seed, T)
    wide = {n: torch.zeros(2, T + 4) for n in O.STEMS}
    for n in O.STEMS:
        wide[n][:, off:off + T] = clip[n] * (0.02 if n == "other" else 1.0)
    dev_wide = {n: v.to(cuda_device) for n, v in wide.items()}
    fe = pkg.MixingFeatureExtractor()
    got = fe.gated_loudness({n: v[:, off:off + T] for n, v in dev_wide.items()})
    ref = B.stems_gated_loudness({n: wide[n][:, off:off + T].numpy() for n in O.STEMS}, 44100)
    assert np.abs(got.cpu().numpy().astype(np.float64) - ref).max() < 1e-3, (got.cpu().numpy(), ref)


@pytest.mark.parametrize("T", [132300, 132301, 200001, 441000, 1323000])
def test_short_term_loudness_iqr_vs_oracle(cuda_device, T):
    """DR_IQR extension (k7_short_iqr): 3 s to 30 s clips, levels that move inside the clip."""
    clip = O.musical_clip(T % 89, T)
    t = torch.arange(T) / 44100.0
    clip["vocals"] = clip["vocals"] * (0.2 + 0.8 * (torch.sin(2 * torch.pi * 0.11 * t) ** 2))     # slow swell
    clip["bass"] = clip["bass"] * torch.where(t < T / 88200.0, 1.0, 0.1)                          # -20 dB step half way
    clip["other"] = torch.cat([clip["other"][:, :T // 3], torch.zeros(2, T - T // 3)], dim=1)     # silent tail: gated
    fe = pkg.MixingFeatureExtractor()
    n0 = pkg.launch_count()
    got = fe.short_term_loudness_iqr({n: v.to(cuda_device) for n, v in clip.items()})
    assert pkg.launch_count() == n0 + 2
    assert got.shape == (5,) and got.dtype == torch.float32
    ref = B.stems_short_term_iqr({n: v.numpy() for n, v in clip.items()}, 44100)
    assert np.abs(got.cpu().numpy().astype(np.float64) - ref).max() < 2e-3, (got.cpu().numpy(), ref)


def test_short_term_iqr_batched_silence_and_limits(cuda_device):
    fe = pkg.MixingFeatureExtractor()
    T = 150000
    clips = [O.musical_clip(60 + i, T) for i in range(3)]
    clips[2]["drums"] = torch.zeros(2, T)
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    got = fe.short_term_loudness_iqr(batch)
    assert got.shape == (3, 5) and float(got[2, 2]) == 0.0                     # no window passes the gate
    for i, c in enumerate(clips):
        one = fe.short_term_loudness_iqr({n: v.to(cuda_device) for n, v in c.items()})
        assert torch.equal(one, got[i])
    with pytest.raises(ValueError, match="3 s"):
        fe.short_term_loudness_iqr({n: torch.zeros(2, 132299, device=cuda_device) for n in O.STEMS})
