"""CPU: the oracle (oracle/mixfe_oracle.py) against the golden vectors recorded from the real
reference (tests/golden/make_golden.py).  This is what pins the oracle on boxes without /root/reference."""
import numpy as np
import pytest
import torch

from oracle import mixfe_oracle as O
from tests._util import golden

G = golden()


def _eq(a: torch.Tensor, key: str, rtol: float = 1e-5, atol: float = 2e-6):
    """Bit-exact on the host that recorded the fixture.  torch's CPU kernels (pocketfft/MKL, vectorised sums)
    round differently per ISA (AVX2 vs AVX-512 hosts differ by a few ULP: 3.4e-6 relative at worst, measured),
    so elsewhere the bound is 10x tighter than the product's 1e-4 tolerance; clamp/gate values (-100, 0, 0.5)
    must still be exact."""
    ref = torch.from_numpy(G[key])
    assert a.shape == ref.shape
    if torch.equal(a, ref):
        return
    if key.startswith("logmel"):
        # log of a power that can sit 1e-7 below its frame's peak: compare the powers, relative to the frame's
        # largest mel band (the criterion of tests/_util.assert_melpower_close, 10x tighter)
        pa, pr = a.double().exp(), ref.double().exp()
        d, tol = (pa - pr).abs(), rtol * pr + 1e-7 * pr.amax(dim=-2, keepdim=True)
        assert bool((d <= tol).all()), f"{key}: mel power worst ratio {float((d / tol).max()):.3g}"
        return
    d = (a.double() - ref.double()).abs()
    tol = rtol * ref.double().abs() + atol
    assert bool((d <= tol).all()), f"{key}: max diff {float(d.max()):.3g}, worst ratio {float((d / tol).max()):.3g}"
    pinned = (ref == -100.0) | (ref == 100.0)
    assert torch.equal(a[pinned], ref[pinned]), f"{key}: a clamped value moved"


@pytest.mark.parametrize("mk,seed,T", [("synthetic_clip", 1234, 66150), ("musical_clip", 1234, 66150),
                                       ("musical_clip", 9, 4097), ("musical_clip", 21, 2048)])
def test_features_bit_exact(mk, seed, T):
    _eq(O.mixing_features(getattr(O, mk)(seed, T)), f"features/{mk}/{seed}/{T}")


def test_features_full_length_clip():
    _eq(O.mixing_features(O.synthetic_clip(1234, 441000)), "features/synthetic_clip/1234/441000")


@pytest.mark.parametrize("nb", [32, 16, 128])
def test_detailed_spectral(nb):
    cfg = O.FrontEndConfig(use_detailed_spectral=True, n_spectral_bins=nb)
    f = O.mixing_features(O.musical_clip(5, 50000), cfg)
    assert f.shape[0] == cfg.feature_dim
    _eq(f, f"features_detailed{nb}/musical_clip/5/50000")


def test_special_inputs():
    T = 44100
    _eq(O.mixing_features({n: torch.zeros(2, T) for n in O.STEMS}), "features/silence")
    g = torch.Generator().manual_seed(1)
    m = torch.randn(T, generator=g) * 0.1
    _eq(O.mixing_features({n: torch.stack([m, m]) for n in O.STEMS}), "features/L_eq_R")
    _eq(O.mixing_features({n: torch.stack([m, -m]) for n in O.STEMS}), "features/L_eq_negR")
    c = {n: torch.zeros(2, T) for n in O.STEMS}
    c["vocals"] = torch.stack([m, 0.5 * m])
    _eq(O.mixing_features(c), "features/only_vocals")


def test_short_clip_raises():
    with pytest.raises(RuntimeError):
        O.mixing_features({n: torch.zeros(2, 400) for n in O.STEMS})


def test_logmel():
    batch = {n: torch.stack([O.musical_clip(s, 30000)[n] for s in (1, 2)]) for n in O.STEMS}
    lm = O.log_mel(batch)
    _eq(lm[:, :, ::4, ::8].contiguous(), "logmel/musical_clip/1,2/30000/sub")
    st = G["logmel/musical_clip/1,2/30000/stats"]
    assert abs(float(lm.mean()) - st[0]) < 1e-6 and float(lm.min()) == st[1] and float(lm.max()) == st[2]
    _eq(O.log_mel({n: v[None] for n, v in O.synthetic_clip(7, 8192).items()}), "logmel/synthetic_clip/7/8192/full")


@pytest.mark.parametrize("cfg", [(2048, 512, 80), (2048, 512, 128)])
def test_other_stft_configs(cfg):
    c = O.FrontEndConfig(44100, *cfg)
    clip = O.musical_clip(4, 20000)
    lm = O.log_mel({n: v[None] for n, v in clip.items()}, c)
    _eq(lm[:, :, ::2, ::4].contiguous(), f"logmel_{cfg[0]}_{cfg[1]}_{cfg[2]}/musical_clip/4/20000/sub")
    _eq(O.mixing_features(clip, c), f"features_{cfg[0]}_{cfg[1]}_{cfg[2]}/musical_clip/4/20000")


@pytest.mark.parametrize("seed,prob", [(1234, 1.0), (99, 0.5), (7, 0.5), (11, 0.5)])
def test_augmentation(seed, prob):
    T = 30000
    clip = O.musical_clip(seed, T)
    torch.manual_seed(seed)
    params = O.draw_clip_aug(prob=prob)
    y = O.apply_clip_aug(clip, params, fft_reverb=True)
    scale = max(float(np.abs(G[f"augment/{seed}/{prob}/{T}/{n}/sub5"]).max()) for n in O.STEMS)
    for n in O.STEMS:
        ref = torch.from_numpy(G[f"augment/{seed}/{prob}/{T}/{n}/sub5"])
        d = float((y[n][:, ::5] - ref).abs().max()) / scale
        # the reference's 22 050-tap fp32 conv1d carries ~5e-6 of its own accumulation noise
        assert d < 2e-5, (n, d)
        e = float((y[n].double() ** 2).mean())
        assert abs(e - G[f"augment/{seed}/{prob}/{T}/{n}/energy"][0]) <= 1e-4 * e + 1e-12


def test_feature_layout_names():
    names = O.feature_layout()
    assert len(names) == 64
    assert names[0] == "bass_rms_L" and names[30:34] == [f"masking_{m}" for m in O.STEMS] and names[49] == "vocals_rms_L"
