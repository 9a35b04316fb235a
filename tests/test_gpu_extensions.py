"""GPU: the two extensions without a reference counterpart (PARITY UNPINNED: arbitrated by oracle/extensions.py, which
restates the definitions in include/mixfe.h) -- weighted masking (K9) and the envelope-follower compressor (K10)."""
import numpy as np
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import extensions as X
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu


def test_band_weights_match_the_oracle():
    fe = pkg.MixingFeatureExtractor()
    w = fe.band_weights()
    assert w.shape == (128,) and float(w.max()) == 1.0
    assert torch.allclose(w.double(), X.a_weighting_band_weights(), rtol=1e-6, atol=1e-12)
    # A-weighting: about -19 dB at 100 Hz relative to 1 kHz, its maximum sits near 2.5 kHz
    top = 2595.0 * np.log10(1.0 + 22050 / 700.0)
    f = 700.0 * (10.0 ** (np.linspace(0, top, 130)[1:-1] / 2595.0) - 1.0)
    assert 2000 < f[int(w.argmax())] < 3200
    k100, k1000 = int(np.abs(f - 100).argmin()), int(np.abs(f - 1000).argmin())
    assert 10 * np.log10(float(w[k100] / w[k1000])) == pytest.approx(-19.1, abs=1.5)


@pytest.mark.parametrize("beta,tau", [(0.0, 1.0), (0.5, 0.25)])
def test_weighted_masking_vs_oracle(cuda_device, beta, tau):
    fe = pkg.MixingFeatureExtractor()
    clips = [O.musical_clip(60 + i, 40000) for i in range(3)]
    clips[1]["bass"] = clips[1]["bass"] * 0.0                       # a silent stem: never active
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    n0 = pkg.launch_count()
    got = fe.weighted_masking(batch, beta=beta, tau=tau)
    assert pkg.launch_count() - n0 == 3 and got.shape == (3, 4)     # K1 (log-mel only) + 2 launches of K9
    for i, c in enumerate(clips):
        ref = X.weighted_masking(c, beta=beta, tau=tau)
        err = float((got[i].cpu().double() - ref).abs().max())
        print(f"clip {i}: weighted masking {got[i].tolist()} max abs err {err:.3g}")
        assert err < 2e-5
    assert float(got[1, 1]) == 0.0
    one = fe.weighted_masking({n: v.to(cuda_device) for n, v in clips[0].items()}, beta=beta, tau=tau)
    assert one.shape == (4,) and torch.equal(one, got[0])


def test_weighted_masking_differs_from_the_unweighted_feature(cuda_device):
    """Loud low-frequency masking counts less under the A-weighting than in the reference's plain mean."""
    fe = pkg.MixingFeatureExtractor()
    clip = {n: v.to(cuda_device) for n, v in O.musical_clip(5, 40000).items()}
    plain = fe.extract_all_features(clip)[30:34]
    weighted = fe.weighted_masking(clip)
    assert float((plain - weighted).abs().max()) > 1e-3
    assert bool(((weighted >= 0) & (weighted <= 1)).all())


@pytest.mark.parametrize("attack_ms,release_ms", [(5.0, 50.0), (0.0, 100.0), (1.0, 0.0), (20.0, 400.0)])
def test_envelope_compressor_vs_oracle(cuda_device, attack_ms, release_ms):
    aug = pkg.AudioAugmenter()
    x = O.musical_clip(77, 30011)["drums"]                          # transients: the detector matters
    n0 = pkg.launch_count()
    got = aug.apply_envelope_compression(x.to(cuda_device), -24, 3, attack_ms, release_ms)
    assert pkg.launch_count() - n0 == 1 and got.shape == x.shape
    ref = X.envelope_compress(x, -24, 3, attack_ms, release_ms)
    err = float((got.cpu().double() - ref).abs().max()) / float(ref.abs().max())
    print(f"attack {attack_ms} ms release {release_ms} ms: rel max err {err:.3g}")
    assert err < 2e-5


def test_envelope_compressor_limits(cuda_device):
    aug = pkg.AudioAugmenter()
    x = O.musical_clip(78, 20000)["vocals"].to(cuda_device)
    # zero time constants: the reference's memoryless compressor (up to its 1e-8 offset on |x|)
    a = aug.apply_envelope_compression(x, -20, 4, 0.0, 0.0)
    b = aug.apply_compression(x, -20, 4)
    assert float((a - b).abs().max()) <= 2e-6 * float(b.abs().max()) + 2e-8
    # a batch: every row has its own detector, carried exactly across the 4096-sample super-blocks
    batch = torch.stack([x, x.flip(-1)])
    y = aug.apply_envelope_compression(batch, -20, 4, 3.0, 120.0)
    assert torch.equal(y[0], aug.apply_envelope_compression(x, -20, 4, 3.0, 120.0))
    with pytest.raises(ValueError):
        aug.apply_envelope_compression(x, -20, 0)
