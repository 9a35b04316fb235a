"""GPU parity of the generic STFT kernel K1G (csrc/k1g_features.cuh): the constructor arguments the reference's
evaluation / style-transfer callers use, MixingFeatureExtractor(44100, 2048, 512, 80)
(inference/validate_retrieval.py:355, src/train_style_transfer.py:523-533), other power-of-two n_fft, and
K1G against K1 on the training configuration.  Same tolerances as test_gpu_features.py."""
import os
import subprocess
import sys

import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O
from tests._util import assert_features_close, assert_logmel_close, assert_melpower_close

pytestmark = pytest.mark.gpu


def _to(clip, dev):
    return {n: v.to(dev) for n, v in clip.items()}


# n_fft 64 .. 4096: every split of log2(n_fft) into register passes of 3 or 4 stages (csrc/fftg.cuh): 3+3, 4+3, 4+4,
# 3+3+3, 4+3+3, 4+4+3, 4+4+4
CONFIGS = [(2048, 512, 80), (512, 128, 64), (4096, 1024, 128), (256, 64, 40), (1024, 512, 96), (2048, 256, 128),
           (128, 32, 24), (64, 16, 12)]


@pytest.mark.parametrize("n_fft,hop,n_mels", CONFIGS)
@pytest.mark.parametrize("T", [44100, 30001])
def test_features_melpower_logmel_vs_oracle(cuda_device, n_fft, hop, n_mels, T):
    cfg = O.FrontEndConfig(n_fft=n_fft, hop_length=hop, n_mels=n_mels)
    clip = O.musical_clip(n_fft + T, T)
    dclip = _to(clip, cuda_device)
    fe = pkg.MixingFeatureExtractor(44100, n_fft, hop, n_mels)
    assert fe.get_feature_dim() == 64
    got = fe.extract_all_features(dclip)
    assert_features_close(got, O.mixing_features(clip, cfg), fe.feature_names(), f"{n_fft}/{hop}/{n_mels} T={T}")
    ofe = O.MelFrontEnd(cfg)
    ref = torch.cat([ofe.mel_power(clip[n][None]) for n in O.STEMS], dim=1)
    mp = fe._plan(cuda_device).debug_mel_power([dclip[n][None] for n in O.STEMS])
    assert mp.shape == ref.shape == (1, 8, n_mels, 1 + T // hop)
    assert_melpower_close(mp, ref, f"{n_fft}/{hop}/{n_mels} T={T}")
    pre = pkg.MelSpectrogramPreprocessor(44100, n_fft, hop, n_mels)
    lm = pre({n: v[None] for n, v in dclip.items()})
    assert_logmel_close(lm, O.log_mel({n: v[None] for n, v in clip.items()}, cfg, dtype=torch.float64),
                        f"{n_fft}/{hop}/{n_mels} T={T}")


def _fuzz_cases():
    """Seeded irregular configurations: hops that do not divide n_fft (odd, prime), hops above n_fft/2, clips barely
    longer than the reflect padding, ragged last tiles.  n_mels stays where every filter has support."""
    import random
    rnd = random.Random(20260)
    cases = []
    for n_fft in (64, 128, 256, 512, 1024, 2048, 4096):
        for _ in range(2):
            hop = rnd.choice([rnd.randrange(max(8, n_fft // 16), n_fft // 2), rnd.randrange(n_fft // 2, n_fft + 1),
                              rnd.choice([97, 101, 250, 441])])
            hop = max(8, min(hop, n_fft))
            n_mels = rnd.choice([8, 16, 24, 40, 64, 80, 96, 128])
            n_mels = min(n_mels, max(8, n_fft // 12))
            T = rnd.choice([n_fft // 2 + 1 + rnd.randrange(0, 50), n_fft + rnd.randrange(1, 3000),
                            8 * hop + rnd.randrange(0, hop), 20000 + rnd.randrange(0, 5000)])
            cases.append((n_fft, hop, n_mels, T))
    return cases


@pytest.mark.parametrize("n_fft,hop,n_mels,T", _fuzz_cases())
def test_irregular_configurations_vs_oracle(cuda_device, n_fft, hop, n_mels, T):
    cfg = O.FrontEndConfig(n_fft=n_fft, hop_length=hop, n_mels=n_mels)
    clip = O.musical_clip(n_fft * 7 + hop, T)
    dclip = _to(clip, cuda_device)
    fe = pkg.MixingFeatureExtractor(44100, n_fft, hop, n_mels)
    what = f"{n_fft}/{hop}/{n_mels} T={T}"
    assert_features_close(fe.extract_all_features(dclip), O.mixing_features(clip, cfg), fe.feature_names(), what)
    ofe = O.MelFrontEnd(cfg)
    ref = torch.cat([ofe.mel_power(clip[n][None]) for n in O.STEMS], dim=1)
    mp = fe._plan(cuda_device).debug_mel_power([dclip[n][None] for n in O.STEMS])
    assert mp.shape == ref.shape == (1, 8, n_mels, 1 + T // hop)
    assert_melpower_close(mp, ref, what)


def test_generic_kernel_serves_non_1024_plans(cuda_device):
    fe = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)
    assert fe._plan(cuda_device).k1_config()["ctas_per_sm"] == 0          # K1G marker
    if os.environ.get("MIXFE_FORCE_GENERIC", "0") in ("", "0"):              # (the knob routes every plan through K1G)
        assert pkg.MixingFeatureExtractor()._plan(cuda_device).k1_config()["ctas_per_sm"] >= 1


def test_batched_2048_matches_per_clip(cuda_device):
    fe = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)
    clips = [O.musical_clip(100 + i, 40000) for i in range(3)]
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    fb, lb = fe.features_and_logmel(batch)
    for i, c in enumerate(clips):
        f1, l1 = fe.features_and_logmel(_to(c, cuda_device))
        assert torch.equal(fb[i], f1) and torch.equal(lb[i], l1[0] if l1.dim() == 4 else l1)


def test_short_clip_raises_for_2048(cuda_device):
    fe = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)
    with pytest.raises(RuntimeError, match="[Pp]adding"):
        fe.extract_all_features({n: torch.zeros(2, 1024, device=cuda_device) for n in O.STEMS})


_FORCED = r"""
import sys, torch
sys.path.insert(0, %r)
import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O
from tests._util import assert_features_close, assert_melpower_close
dev = torch.device("cuda", 0)
fe = pkg.MixingFeatureExtractor()
assert fe._plan(dev).k1_config()["ctas_per_sm"] == 0
for T in (2048, 30001, 66150):
    clip = O.musical_clip(T, T)
    d = {n: v.to(dev) for n, v in clip.items()}
    assert_features_close(fe.extract_all_features(d), O.mixing_features(clip), fe.feature_names(), "forced generic")
    ofe = O.MelFrontEnd(O.FrontEndConfig())
    ref = torch.cat([ofe.mel_power(clip[n][None]) for n in O.STEMS], dim=1)
    assert_melpower_close(fe._plan(dev).debug_mel_power([d[n][None] for n in O.STEMS]), ref, "forced generic")
print("forced-generic ok")
"""


def test_generic_kernel_on_training_config(cuda_device):
    """MIXFE_FORCE_GENERIC=1 routes (1024, 256, 128) through K1G: same oracle, same tolerances as K1."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, MIXFE_FORCE_GENERIC="1")
    res = subprocess.run([sys.executable, "-c", _FORCED % root], env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "forced-generic ok" in res.stdout, res.stdout + res.stderr


@pytest.mark.parametrize("n_fft,hop,n_mels,nb", [(2048, 512, 80, 32), (2048, 512, 80, 80), (512, 128, 64, 1), (512, 128, 64, 2),
                                                 (1024, 256, 96, 33), (1024, 300, 128, 127), (256, 64, 40, 7)])
def test_detailed_spectral_mode_on_other_configurations(cuda_device, n_fft, hop, n_mels, nb):
    """use_detailed_spectral=True (src/mixing_utils.py:200-224) with bin counts below, at and just under n_mels, on the
    generic kernel and on K1 with a non-default filter count."""
    cfg = O.FrontEndConfig(n_fft=n_fft, hop_length=hop, n_mels=n_mels, use_detailed_spectral=True, n_spectral_bins=nb)
    T = 21013
    clip = O.musical_clip(n_fft + nb, T)
    fe = pkg.MixingFeatureExtractor(44100, n_fft, hop, n_mels, use_detailed_spectral=True, n_spectral_bins=nb)
    assert fe.get_feature_dim() == cfg.feature_dim
    got = fe.extract_all_features(_to(clip, cuda_device))
    assert_features_close(got, O.mixing_features(clip, cfg), fe.feature_names(), f"{n_fft}/{hop}/{n_mels} detailed {nb}")


def test_second_configuration_from_one_read_of_the_audio(cuda_device):
    """features_and_logmel(..., second=(2048, 512, 80)): the extractor's own configuration plus the frozen encoder's,
    chunked so that the second kernel reads the samples out of L2.  Same numbers as two
    separate calls, bit for bit, and the reference's golden log-mel of the second configuration."""
    import numpy as np
    from tests._util import golden
    G = golden()
    clips = [O.musical_clip(4, 20000)] + [O.musical_clip(90 + i, 20000) for i in range(4)]
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    fe = pkg.MixingFeatureExtractor()
    f, lm, lm2 = fe.features_and_logmel(batch, second=(2048, 512, 80), chunk_clips=2)
    f0, lm0 = fe.features_and_logmel(batch)
    assert torch.equal(f, f0) and torch.equal(lm, lm0)
    sep = pkg.MelSpectrogramPreprocessor(44100, 2048, 512, 80)(batch)
    assert lm2.shape == sep.shape == (5, 8, 80, 1 + 20000 // 512) and torch.equal(lm2, sep)
    ref_sub = torch.from_numpy(G["logmel_2048_512_80/musical_clip/4/20000/sub"])
    d = (lm2[0:1, :, ::2, ::4].cpu() - ref_sub).abs()
    assert float(d.mean()) < 2e-5
    cfg = O.FrontEndConfig(n_fft=2048, hop_length=512, n_mels=80)
    assert_logmel_close(lm2[0:1], O.log_mel({n: v[None] for n, v in clips[0].items()}, cfg, dtype=torch.float64), "second config")
    one = fe.features_and_logmel({n: v.to(cuda_device) for n, v in clips[1].items()}, second=(2048, 512, 80))
    assert one[0].shape == (64,) and torch.equal(one[0], f[1]) and torch.equal(one[2][0], lm2[1])


@pytest.mark.parametrize("hop,n_mels,T", [(512, 80, 1025), (512, 80, 2047), (512, 80, 2048), (512, 80, 4097), (512, 80, 10240 + 3),
                                          (512, 128, 30000), (256, 128, 9001), (1024, 64, 20000), (2048, 40, 33333), (512, 17, 8000),
                                          (100, 80, 7000)])
def test_k1e_lengths_hops_and_unaligned_views(cuda_device, hop, n_mels, T):
    """K1E (log-mel of n_fft = 2048 plans; two warps per frame): clip lengths around its 4-frame tiles and the 1024-sample
    reflect halo, hops that are multiples of 4, ragged filter counts, and views that start 1..3 floats into a wider
    tensor (cooperative staging instead of TMA).  Against the fp64 oracle."""
    cfg = O.FrontEndConfig(n_fft=2048, hop_length=hop, n_mels=n_mels)
    clips = [O.musical_clip(500 + T % 97 + i, T) for i in range(2)]
    off = 1 + T % 3
    wide = torch.zeros(2, 8, T + 8, device=cuda_device)
    for i, c in enumerate(clips):
        for k, n in enumerate(O.STEMS):
            wide[i, 2 * k:2 * k + 2, off:off + T] = c[n].to(cuda_device)
    pre = pkg.MelSpectrogramPreprocessor(44100, 2048, hop, n_mels)
    views = {n: wide[:, 2 * k:2 * k + 2, off:off + T] for k, n in enumerate(O.STEMS)}
    n0 = pkg.launch_count()
    lm_u = pre(views)
    assert pkg.launch_count() == n0 + 1
    lm_a = pre({n: v.contiguous() for n, v in views.items()})         # aligned copy: the TMA path
    assert torch.equal(lm_u, lm_a)
    assert lm_u.shape == (2, 8, n_mels, 1 + T // hop)
    for i, c in enumerate(clips):
        ref = O.log_mel({n: v[None] for n, v in c.items()}, cfg, dtype=torch.float64)
        assert_logmel_close(lm_u[i:i + 1], ref, f"K1E hop={hop} n_mels={n_mels} T={T} clip {i}")


@pytest.mark.parametrize("hop,n_mels,T", [(512, 80, 1025), (512, 80, 2048), (512, 80, 4097), (512, 80, 10240 + 3), (512, 128, 30000),
                                          (256, 96, 9001), (1024, 64, 20000), (520, 80, 12345), (96, 40, 7000), (2048, 40, 33333)])
def test_k1e_features_lengths_hops_and_unaligned_views(cuda_device, hop, n_mels, T):
    """K1E with statistics (mixing features of n_fft = 2048 plans whose hop is a multiple of 8): clip lengths around the
    4-frame tiles, hops with (32 | hop: TMA statistics rows) and without (global role loads), ragged filter counts, and
    views that start 1..3 floats into a wider tensor.  Features against the oracle; aligned and unaligned inputs agree
    bit for bit; K1 + K2 = two launches."""
    cfg = O.FrontEndConfig(n_fft=2048, hop_length=hop, n_mels=n_mels)
    clips = [O.musical_clip(900 + T % 89 + i, T) for i in range(2)]
    off = 1 + T % 3
    wide = torch.zeros(2, 8, T + 8, device=cuda_device)
    for i, c in enumerate(clips):
        for k, n in enumerate(O.STEMS):
            wide[i, 2 * k:2 * k + 2, off:off + T] = c[n].to(cuda_device)
    fe = pkg.MixingFeatureExtractor(44100, 2048, hop, n_mels)
    views = {n: wide[:, 2 * k:2 * k + 2, off:off + T] for k, n in enumerate(O.STEMS)}
    n0 = pkg.launch_count()
    f_u, lm_u = fe.features_and_logmel(views)
    assert pkg.launch_count() == n0 + 2
    f_a, lm_a = fe.features_and_logmel({n: v.contiguous() for n, v in views.items()})
    assert torch.equal(lm_u, lm_a)
    assert torch.allclose(f_u, f_a, rtol=1e-5, atol=1e-6)              # role loads and TMA rows sum in different orders
    f_only = fe.extract_all_features(views)
    assert torch.equal(f_only, f_u)
    for i, c in enumerate(clips):
        assert_features_close(f_u[i], O.mixing_features(c, cfg), fe.feature_names(), f"K1E hop={hop} n_mels={n_mels} T={T} clip {i}")
        ref = O.log_mel({n: v[None] for n, v in c.items()}, cfg, dtype=torch.float64)
        assert_logmel_close(lm_u[i:i + 1], ref, f"K1E hop={hop} n_mels={n_mels} T={T} clip {i}")
