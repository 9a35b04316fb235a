"""GPU: the four single-effect methods of AudioAugmenter (src/mixing_utils.py:421-479) through mixfe_effects /
mixfe_reverb, against round-2 golden vectors of the real reference and against the oracle, and the song_A
known-answer vector (BASELINE config 1) through the feature kernels."""
import os

import numpy as np
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O
from tests._util import assert_features_close, assert_logmel_close, golden

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
G2 = np.load(os.path.join(HERE, "golden", "reference_outputs_r2.npz"))
X = O.musical_clip(77, 30000)["drums"]


def _rel(got, ref):
    ref = torch.as_tensor(ref)
    return float((got.cpu() - ref).abs().max()) / float(ref.abs().max())


@pytest.mark.parametrize("seed,decay", [(5, 0.5), (6, 0.25), (7, 0.1)])
def test_apply_reverb_vs_reference(cuda_device, seed, decay):
    aug = pkg.AudioAugmenter()
    n0 = pkg.launch_count()
    torch.manual_seed(seed)
    got = aug.apply_reverb(X, decay)                 # CPU tensor in: impulse drawn on the CPU generator like the reference
    assert pkg.launch_count() - n0 == 6              # parameter fetch + the reverb's five launches: no K3, no redistribution
    assert got.device.type == "cpu" and got.shape == X.shape
    assert _rel(got, G2[f"apply_reverb/{seed}/{decay}"]) < 2e-5
    torch.manual_seed(seed)
    h = O.draw_impulse(decay)
    got2 = aug.apply_reverb(X.to(cuda_device), decay, impulse=h)
    assert got2.device == cuda_device
    assert _rel(got2, O.reverb_standalone(X, h)) < 5e-6
    assert torch.equal(got2.cpu(), got)


def test_apply_reverb_batched_mono_and_wide(cuda_device):
    aug = pkg.AudioAugmenter()
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3, 2, 25001, generator=g) * 0.1
    h = torch.randn(3, 1500, generator=g) * 0.05
    got = aug.apply_reverb(x.to(cuda_device), impulse=h)
    for b in range(3):
        assert _rel(got[b], O.reverb_standalone(x[b], h[b])) < 5e-6
    mono = aug.apply_reverb(x[0, :1].to(cuda_device), impulse=h[0])
    assert mono.shape == (1, 25001) and _rel(mono, O.reverb_standalone(x[0, :1], h[0])) < 5e-6
    wide = torch.cat([x[0], x[1], x[2][:1]])         # 5 channels, one impulse (what the reference's loop does)
    gw = aug.apply_reverb(wide.to(cuda_device), impulse=h[1])
    assert gw.shape == wide.shape and _rel(gw, O.reverb_standalone(wide, h[1])) < 5e-6


@pytest.mark.parametrize("thr,ratio", [(-20, 4), (-30, 2.5), (-12, 8), (-6.5, 3)])
def test_apply_compression_arguments(cuda_device, thr, ratio):
    aug = pkg.AudioAugmenter()
    n0 = pkg.launch_count()
    got = aug.apply_compression(X.to(cuda_device), thr, ratio)
    assert pkg.launch_count() - n0 == 2              # parameter fetch + K3
    ref = torch.from_numpy(G2[f"apply_compression/{thr}/{ratio}"])
    err = _rel(got, ref)
    print(f"compression thr={thr} ratio={ratio}: rel max err {err:.3g}")
    assert err < 5e-6
    # the mask decision (|x| + 1e-8 above / below the knee) away from the knee is the reference's
    a = X.abs() + 1e-8
    far = (20 * torch.log10(a) - thr).abs() > 1e-3
    mine = (got.cpu().abs() < a * (1 - 1e-5)) | (got.cpu().abs() > a * (1 + 1e-5))
    refm = (20 * torch.log10(a)) > thr
    assert torch.equal(mine[far & (a > 1e-6)], refm[far & (a > 1e-6)]) or thr > 0
    with pytest.raises(ValueError):
        aug.apply_compression(X.to(cuda_device), -20, 0)


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_tilt_and_bandwidth_vs_reference(cuda_device, seed):
    aug = pkg.AudioAugmenter()
    torch.manual_seed(seed)
    e1 = _rel(aug.apply_spectral_tilt(X.to(cuda_device)), G2[f"apply_spectral_tilt/{seed}"])
    torch.manual_seed(seed)
    e2 = _rel(aug.apply_bandwidth_limit(X.to(cuda_device)), G2[f"apply_bandwidth_limit/{seed}"])
    print(f"seed {seed}: tilt {e1:.3g} band-limit {e2:.3g}")
    assert e1 < 2e-5 and e2 < 2e-5


def test_effects_batched_and_strided(cuda_device):
    aug = pkg.AudioAugmenter()
    g = torch.Generator().manual_seed(9)
    x = torch.randn(4, 3, 12345, generator=g) * 0.2                    # 3 channels per clip, odd length
    wide = torch.zeros(4, 3, 12350, device=cuda_device)
    wide[:, :, 3:12348] = x.to(cuda_device)
    got = aug.apply_compression(wide[:, :, 3:12348], -18, 3)
    ref = O.compress_static(x, -18, 3)
    assert got.shape == x.shape and _rel(got, ref) < 5e-6
    assert float(wide[:, :, :3].abs().max()) == 0.0 and float(wide[:, :, 12348:].abs().max()) == 0.0


def test_augment_plan_has_no_feature_tables(cuda_device):
    aug = pkg.AudioAugmenter()
    plan = aug._plan(cuda_device)
    with pytest.raises(ValueError, match="augmentation-only"):
        plan.features_logmel([torch.zeros(1, 2, 4096, device=cuda_device)] * 4)


def test_highpass_of_low_frequency_input_keeps_its_accuracy(cuda_device):
    """A 60 Hz tone through the 2 kHz high-pass: the output is ~60 dB under the input; the lane-local fp32
    arithmetic must not let the error scale with the INPUT (k3_augment.cuh header)."""
    T = 200000
    t = torch.arange(T, dtype=torch.float64) / 44100
    g = torch.Generator().manual_seed(4)
    x = (0.5 * torch.sin(2 * np.pi * 60 * t) + 1e-3 * torch.randn(T, generator=g, dtype=torch.float64)).float()
    x = torch.stack([x, -x])
    aug = pkg.AudioAugmenter()
    for seed in range(40):
        torch.manual_seed(seed)
        if torch.rand(1) < 0.5:
            break
    torch.manual_seed(seed)
    got = aug.apply_spectral_tilt(x.to(cuda_device))
    ref = O.spectral_tilt(x, "high")
    err = float((got.cpu() - ref)[:, 2000:].abs().max()) / float(ref[:, 2000:].abs().max())
    print(f"high-pass of a 60 Hz tone: rel max err {err:.3g}, output peak {float(ref[:, 2000:].abs().max()):.3g}")
    assert err < 1e-4


def test_song_A_features_and_logmel(cuda_device):
    """BASELINE config 1 on the GPU: the reference's own asset, features and log-mel against the values recorded from
    the real reference and against the oracle."""
    G = golden()
    pcm = np.load(os.path.join(HERE, "golden", "song_A_int16.npz"))["pcm"]
    clip = O.song_a_clip(pcm)
    fe = pkg.MixingFeatureExtractor()
    feats, lm = fe.features_and_logmel({n: v.to(cuda_device) for n, v in clip.items()})
    worst = assert_features_close(feats, torch.from_numpy(G["features/song_A"]), fe.feature_names(), "song_A")
    print(f"song_A: max abs feature error vs the reference {worst:.3g}")
    st = G["logmel/song_A/stats"]
    assert lm.shape == (1, 8, 128, 1723)
    assert abs(float(lm.double().mean()) - st[0]) < 2e-5
    assert float(lm.min()) == pytest.approx(st[1], abs=1e-6) and float(lm.max()) == pytest.approx(st[2], abs=2e-5)
    assert float(lm[0, 0, 64, 861]) == pytest.approx(st[3], abs=5e-5)
    ref64 = O.log_mel({n: v[None] for n, v in clip.items()}, dtype=torch.float64)
    assert_logmel_close(lm, ref64, "song_A log-mel")
