"""GPU parity: libmixfe (through the Python drop-ins -> ctypes -> C ABI) against the CPU oracle and the
golden vectors of the real reference.  Tolerances are BASELINE.md section 5:
features |d| <= 1e-4|ref| + 1e-5; mel power |d| <= 1e-4|ref| + 1e-6 * frame max; log-mel max-abs
reported (the fp32 oracle itself is only good to ~1e-3 there)."""
import numpy as np
import os

import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O
from tests._util import assert_features_close, assert_logmel_close, assert_melpower_close, golden

pytestmark = pytest.mark.gpu
G = golden()


def _to(clip, dev):
    return {n: v.to(dev) for n, v in clip.items()}


@pytest.fixture(scope="module")
def fe(cuda_device):
    return pkg.MixingFeatureExtractor()


@pytest.mark.parametrize("mk,seed,T", [("synthetic_clip", 1234, 66150), ("musical_clip", 1234, 66150),
                                       ("musical_clip", 9, 4097), ("musical_clip", 21, 2048),
                                       ("synthetic_clip", 1234, 441000)])
def test_features_vs_reference_golden(fe, cuda_device, mk, seed, T):
    clip = getattr(O, mk)(seed, T)
    n0 = pkg.launch_count()
    got = fe.extract_all_features(_to(clip, cuda_device))
    assert pkg.launch_count() == n0 + 2                          # K1 + K2, nothing else
    assert got.shape == (64,) and got.device == cuda_device and got.dtype == torch.float32
    ref = torch.from_numpy(G[f"features/{mk}/{seed}/{T}"])
    assert_features_close(got, ref, fe.feature_names(), f"{mk}/{seed}/{T}")


@pytest.mark.parametrize("T", [513, 1000, 1024, 1279, 1280, 2047, 2048, 2049, 4096, 8191, 30000, 44100])
def test_features_and_melpower_vs_oracle_ragged_lengths(fe, cuda_device, T):
    """Every tile shape: clips shorter than one tile, T a multiple of hop (last frame starts at T), odd T
    (misaligned right channel -> scalar staging path)."""
    clip = O.musical_clip(T, T)
    dclip = _to(clip, cuda_device)
    assert_features_close(fe.extract_all_features(dclip), O.mixing_features(clip), fe.feature_names(), f"T={T}")
    ofe = O.MelFrontEnd(O.FrontEndConfig())
    ref = torch.cat([ofe.mel_power(clip[n][None]) for n in O.STEMS], dim=1)
    got = fe._plan(cuda_device).debug_mel_power([dclip[n][None] for n in O.STEMS])
    assert got.shape == ref.shape == (1, 8, 128, 1 + T // 256)
    assert_melpower_close(got, ref, f"T={T}")


def test_short_clip_raises_like_torch_stft(fe, cuda_device):
    for T in (400, 512):
        with pytest.raises(RuntimeError, match="[Pp]adding"):
            fe.extract_all_features({n: torch.zeros(2, T, device=cuda_device) for n in O.STEMS})


def test_missing_stem_raises_keyerror(fe, cuda_device):
    clip = _to(O.synthetic_clip(0, 4096), cuda_device)
    del clip["drums"]
    with pytest.raises(KeyError):
        fe.extract_all_features(clip)


def test_special_inputs(fe, cuda_device):
    T = 44100
    names = fe.feature_names()
    z = {n: torch.zeros(2, T, device=cuda_device) for n in O.STEMS}
    assert_features_close(fe.extract_all_features(z), torch.from_numpy(G["features/silence"]), names, "silence")
    g = torch.Generator().manual_seed(1)
    m = (torch.randn(T, generator=g) * 0.1).to(cuda_device)
    for key, pair in (("L_eq_R", torch.stack([m, m])), ("L_eq_negR", torch.stack([m, -m]))):
        got = fe.extract_all_features({n: pair.clone() for n in O.STEMS})
        assert_features_close(got, torch.from_numpy(G[f"features/{key}"]), names, key)
    c = {n: torch.zeros(2, T, device=cuda_device) for n in O.STEMS}
    c["vocals"] = torch.stack([m, 0.5 * m])
    assert_features_close(fe.extract_all_features(c), torch.from_numpy(G["features/only_vocals"]), names, "only vocals")


def test_nan_is_scrubbed_and_values_clamped(fe, cuda_device):
    clip = _to(O.synthetic_clip(5, 8192), cuda_device)
    clip["bass"] = clip["bass"].clone()
    clip["bass"][0, 100] = float("nan")
    f = fe.extract_all_features(clip)
    assert bool(torch.isfinite(f).all()) and float(f.abs().max()) <= 100.0


@pytest.mark.parametrize("nb", [32, 16, 128])
def test_detailed_spectral_mode(cuda_device, nb):
    fd = pkg.MixingFeatureExtractor(use_detailed_spectral=True, n_spectral_bins=nb)
    clip = O.musical_clip(5, 50000)
    got = fd.extract_all_features(_to(clip, cuda_device))
    assert got.shape[0] == fd.get_feature_dim()
    assert_features_close(got, torch.from_numpy(G[f"features_detailed{nb}/musical_clip/5/50000"]), fd.feature_names(), f"detailed {nb}")


def test_detailed_bins_above_mels_raises(cuda_device):
    fd = pkg.MixingFeatureExtractor(use_detailed_spectral=True, n_spectral_bins=200)
    with pytest.raises(RuntimeError):
        fd.extract_all_features(_to(O.synthetic_clip(0, 4096), cuda_device))


def test_logmel_module_vs_golden_and_oracle(cuda_device):
    mod = pkg.MelSpectrogramPreprocessor(44100, 1024, 256, 128).to(cuda_device)
    batch = {n: torch.stack([O.musical_clip(s, 30000)[n] for s in (1, 2)]) for n in O.STEMS}
    lm = mod(_to(batch, cuda_device))
    assert lm.shape == (2, 8, 128, 118) and lm.is_contiguous()
    ref_sub = torch.from_numpy(G["logmel/musical_clip/1,2/30000/sub"])
    d = (lm[:, :, ::4, ::8].cpu() - ref_sub).abs()
    # quiet bins of the fp32 reference carry ~1e-3 of the reference's own FFT noise
    assert float(d.max()) < 5e-3 and float(d.mean()) < 2e-5, (float(d.max()), float(d.mean()))
    # and exact agreement wherever the bin is not quiet
    ref_full = O.log_mel(batch)
    loud = ref_full > ref_full.amax(dim=2, keepdim=True) - 9.0          # within ~40 dB of the frame maximum
    assert float((lm.cpu() - ref_full)[loud].abs().max()) < 2e-4
    assert float(lm[:, :, 0, :].max()) == pytest.approx(np.log(1e-10), abs=1e-5)      # mel filter 0 is all-zero
    full = torch.from_numpy(G["logmel/synthetic_clip/7/8192/full"])
    got = mod(_to({n: v[None] for n, v in O.synthetic_clip(7, 8192).items()}, cuda_device)).cpu()
    assert float((got - full).abs().max()) < 2e-4


def test_mel_module_checkpoint_roundtrip_and_fb_override(cuda_device):
    mod = pkg.MelSpectrogramPreprocessor().to(cuda_device)
    clip = _to({n: v[None] for n, v in O.synthetic_clip(3, 8192).items()}, cuda_device)
    a = mod(clip)
    sd = {k: v.clone() for k, v in mod.state_dict().items()}
    sd["mel_transform.mel_scale.fb"] *= 2.0
    mod.load_state_dict(sd, strict=True)
    b = mod(clip)
    # doubling the filterbank doubles the mel power: log shifts by ln 2 wherever mel >> 1e-10
    mask = a > -10
    assert float(((b - a)[mask] - np.log(2.0)).abs().max()) < 1e-4


def test_batched_equals_per_clip_and_is_deterministic(fe, cuda_device):
    clips = [O.musical_clip(s, 20000) for s in (1, 2, 3)]
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    f, lm = fe.features_and_logmel(batch)
    f2, lm2 = fe.features_and_logmel(batch)
    assert torch.equal(f, f2) and torch.equal(lm, lm2)                       # no float atomics anywhere
    assert f.shape == (3, 64) and lm.shape == (3, 8, 128, 79)
    for i, c in enumerate(clips):
        fi, li = fe.features_and_logmel(_to(c, cuda_device))
        assert torch.equal(fi, f[i]) and torch.equal(li[0], lm[i])


def test_strided_views_and_cpu_inputs(fe, cuda_device):
    """Slices of a (B,8,T) tensor (src/train_style_transfer.py:196,208) and crops (src/data.py:282)."""
    g = torch.Generator().manual_seed(11)
    big = (torch.rand(2, 8, 9000, generator=g) * 2 - 1) * 0.3
    dev_big = big.to(cuda_device)
    views = {n: dev_big[:, 2 * i:2 * i + 2, 500:8692] for i, n in enumerate(O.STEMS)}
    f = fe.extract_all_features(views)
    for b in range(2):
        clip = {n: big[b, 2 * i:2 * i + 2, 500:8692].contiguous() for i, n in enumerate(O.STEMS)}
        assert_features_close(f[b], O.mixing_features(clip), fe.feature_names(), f"view clip {b}")
    cpu_clip = O.synthetic_clip(4, 6000)
    out = fe.extract_all_features(cpu_clip)                                   # H2D, GPU compute, D2H
    assert out.device.type == "cpu"
    assert_features_close(out, O.mixing_features(cpu_clip), fe.feature_names(), "cpu input")


def test_unaligned_rows_take_the_cooperative_staging_path(fe, cuda_device):
    """Rows that are not 16-byte aligned cannot be staged by TMA bulk copies nor read with 128-bit loads:
    interior tiles then go through the cooperative fallback.  Same numbers either way."""
    g = torch.Generator().manual_seed(12)
    big = (torch.rand(2, 8, 12000, generator=g) * 2 - 1) * 0.3
    dev_big = big.to(cuda_device)
    for off, T in ((501, 9001), (3, 11997), (500, 9000)):
        views = {n: dev_big[:, 2 * i:2 * i + 2, off:off + T] for i, n in enumerate(O.STEMS)}
        f, lm = fe.features_and_logmel(views)
        for b in range(2):
            clip = {n: big[b, 2 * i:2 * i + 2, off:off + T].contiguous() for i, n in enumerate(O.STEMS)}
            assert_features_close(f[b], O.mixing_features(clip), fe.feature_names(), f"offset {off} clip {b}")
            assert_logmel_close(lm[b:b + 1], O.log_mel({n: v[None] for n, v in clip.items()}, dtype=torch.float64),
                                f"offset {off} clip {b}")
        # and bit-identical to the aligned (TMA) path on a contiguous copy of the same samples
        f2, lm2 = fe.features_and_logmel({n: v.contiguous() for n, v in views.items()})
        assert torch.equal(lm, lm2)
        assert torch.allclose(f, f2, rtol=1e-5, atol=1e-6)           # statistics sum in a different order


def test_thirty_second_clips_of_the_offline_sweep(fe, cuda_device):
    """BASELINE config 4 shape: 30 s clips (T = 1 323 000, F = 5168); clip 0 against the oracle."""
    B, T = 3, 1323000
    g = torch.Generator(device=cuda_device).manual_seed(4096)
    batch = {n: (torch.rand(B, 2, T, device=cuda_device, generator=g) * 2 - 1) * 0.2 for n in O.STEMS}
    f, lm = fe.features_and_logmel(batch)
    assert lm.shape == (B, 8, 128, 5168) and f.shape == (B, 64)
    clip0 = {n: batch[n][0].cpu() for n in O.STEMS}
    assert_features_close(f[0], O.mixing_features(clip0), fe.feature_names(), "30 s clip 0")
    assert_logmel_close(lm[0:1], O.log_mel({n: v[None] for n, v in clip0.items()}, dtype=torch.float64), "30 s clip 0")
    if os.environ.get("MIXFE_FORCE_GENERIC", "0") in ("", "0"):       # (the knob routes the plan through K1G)
        assert fe._plan(cuda_device).k1_config()["ctas_per_sm"] == 2     # the shared-memory budget keeps 2 CTAs / SM


def test_full_size_properties(fe, cuda_device):
    """BASELINE config 2 shape (a slice of it): properties that need no oracle run.
    Gain linearity: scaling every stem by 2 adds ln 4 to the log-mel (where mel >> 1e-10), leaves the
    stereo / masking-free ratios alone and adds 20 log10 2 to the loudness features."""
    B, T = 6, 441000
    g = torch.Generator(device=cuda_device).manual_seed(20261104)
    batch = {n: (torch.rand(B, 2, T, device=cuda_device, generator=g) * 2 - 1) * 0.25 for n in O.STEMS}
    f1, l1 = fe.features_and_logmel(batch)
    f2, l2 = fe.features_and_logmel({n: 2 * v for n, v in batch.items()})
    assert l1.shape == (B, 8, 128, 1723)
    body = l1[:, :, 1:, :]
    assert float(((l2[:, :, 1:, :] - body) - np.log(4.0)).abs().max()) < 1e-4
    names = fe.feature_names()
    for i, nme in enumerate(names):
        d = (f2[:, i] - f1[:, i]).abs().max().item()
        if "rms" in nme:
            assert torch.allclose(f2[:, i], 2 * f1[:, i], rtol=1e-5)
        elif nme.endswith(("loudness", "loudness_dup", "spectral_low", "spectral_mid", "spectral_high")):
            pass                                                     # shifted by a constant; checked below
        elif "masking" in nme or "flatness" in nme or "tilt" in nme:
            pass                                                     # not scale invariant (mel band 0 is pinned at -100 dB)
        else:
            assert d < 2e-4, (nme, d)
    i = names.index("vocals_loudness")
    assert torch.allclose(f2[:, i] - f1[:, i], torch.full((B,), 20 * np.log10(2.0), device=cuda_device), atol=1e-4)
    # clip 0 alone against the oracle (one full-length clip keeps the CPU side to a second)
    clip0 = {n: batch[n][0].cpu() for n in O.STEMS}
    assert_features_close(f1[0], O.mixing_features(clip0), names, "full-size clip 0")


def test_smoke_entry(cuda_device):
    import __graft_entry__ as ge
    ge.smoke()


@pytest.mark.parametrize("sr,n_mels", [(8000, 128), (16000, 128), (16000, 64), (22050, 128), (32000, 80), (48000, 128),
                                       (96000, 128), (44100, 8), (44100, 1), (11025, 33)])
def test_other_sample_rates_and_filter_counts(cuda_device, sr, n_mels):
    """The mel walk's tables (groups of 8 filters, start residues, empty filters) depend on the filterbank's shape:
    low sample rates give filters narrower than a bin (some empty), n_mels not a multiple of 8 gives a ragged last
    group.  Features, mel power and log-mel against the oracle."""
    cfg = O.FrontEndConfig(sample_rate=sr, n_mels=n_mels)
    T = 30011
    clip = O.musical_clip(sr % 1000 + n_mels, T, sample_rate=sr)
    dclip = _to(clip, cuda_device)
    fe = pkg.MixingFeatureExtractor(sr, 1024, 256, n_mels)
    what = f"sr={sr} n_mels={n_mels}"
    assert_features_close(fe.extract_all_features(dclip), O.mixing_features(clip, cfg), fe.feature_names(), what)
    ofe = O.MelFrontEnd(cfg)
    ref = torch.cat([ofe.mel_power(clip[n][None]) for n in O.STEMS], dim=1)
    mp = fe._plan(cuda_device).debug_mel_power([dclip[n][None] for n in O.STEMS])
    assert_melpower_close(mp, ref, what)
    lm = pkg.MelSpectrogramPreprocessor(sr, 1024, 256, n_mels)({n: v[None] for n, v in dclip.items()})
    assert_logmel_close(lm, O.log_mel({n: v[None] for n, v in clip.items()}, cfg, dtype=torch.float64), what)


@pytest.mark.parametrize("Tw,T", [(12001, 9000), (12002, 9000), (12003, 11000), (12004, 9000)])
def test_aligned_base_with_unaligned_row_strides(fe, cuda_device, Tw, T):
    """The first row is 16-byte aligned but the channel / clip strides are not multiples of 4 floats (Tw = 12001..12003):
    later rows are misaligned, so the host must not pick the TMA staging for the launch.  Tw = 12004 is the aligned
    control.  Same numbers as a contiguous copy, bit for bit in the log-mel."""
    g = torch.Generator().manual_seed(Tw)
    big = (torch.rand(3, 8, Tw, generator=g) * 2 - 1) * 0.3
    dev_big = big.to(cuda_device)
    views = {n: dev_big[:, 2 * i:2 * i + 2, :T] for i, n in enumerate(O.STEMS)}
    f, lm = fe.features_and_logmel(views)
    f2, lm2 = fe.features_and_logmel({n: v.contiguous() for n, v in views.items()})
    assert torch.equal(lm, lm2)
    assert torch.allclose(f, f2, rtol=1e-5, atol=1e-6)
    clip = {n: big[2, 2 * i:2 * i + 2, :T].contiguous() for i, n in enumerate(O.STEMS)}
    assert_features_close(f[2], O.mixing_features(clip), fe.feature_names(), f"Tw={Tw} clip 2")
