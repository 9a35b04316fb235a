"""GPU parity of K6 (csrc/k6_logmel_bwd.cuh): d loss / d audio through MelSpectrogramPreprocessor, against torch
autograd through the CPU oracle's log-mel (the reference's path: src/model.py:41-67 under autograd, as the
style-transfer loop uses it, src/train_style_transfer.py:206-215).

Tolerance.  The gradient carries 1 / (mel + 1e-10), so fp32 itself is far from exact on quiet bins; the test is
self-calibrating: err(ours vs fp64 oracle) <= max(4 x err(fp32 oracle vs fp64 oracle), 1e-4 x max|grad|), all
in the max norm over a stem."""
import os

import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu


def _oracle_grad(clip, cfg, gout, dtype):
    leaves = {n: clip[n][None].to(dtype).clone().requires_grad_(True) for n in O.STEMS}
    lm = O.log_mel(leaves, cfg, dtype=dtype)
    (lm * gout.to(dtype)).sum().backward()
    return {n: leaves[n].grad[0].double() for n in O.STEMS}


def _noisy_clip(seed, T):
    clip = O.musical_clip(seed, T)
    g = torch.Generator().manual_seed(seed + 7)
    return {n: v + 0.01 * torch.randn(v.shape, generator=g) for n, v in clip.items()}   # keeps every mel bin audible


@pytest.mark.parametrize("n_fft,hop,n_mels,T", [(1024, 256, 128, 20000), (2048, 512, 80, 30001), (512, 128, 64, 9000),
                                                (4096, 1024, 128, 40000), (1024, 256, 128, 1500), (256, 64, 40, 5000),
                                                (128, 32, 24, 3000), (64, 16, 12, 2000),
                                                # hops that divide nothing, hop > n_fft/2, one-frame clips, ragged tiles
                                                (1024, 300, 40, 5810), (1024, 703, 64, 561), (2048, 1087, 96, 9343),
                                                (512, 434, 8, 3700), (256, 93, 16, 2111), (4096, 450, 96, 4323)])
def test_backward_matches_autograd_through_oracle(cuda_device, n_fft, hop, n_mels, T):
    cfg = O.FrontEndConfig(n_fft=n_fft, hop_length=hop, n_mels=n_mels)
    clip = _noisy_clip(n_fft + T, T)
    F = cfg.n_frames(T)
    gout = torch.randn(1, 8, n_mels, F, generator=torch.Generator().manual_seed(3))
    ref64 = _oracle_grad(clip, cfg, gout, torch.float64)
    ref32 = _oracle_grad(clip, cfg, gout, torch.float32)

    pre = pkg.MelSpectrogramPreprocessor(44100, n_fft, hop, n_mels)
    leaves = {n: clip[n][None].to(cuda_device).requires_grad_(True) for n in O.STEMS}
    n0 = pkg.launch_count()
    lm = pre(leaves)
    assert lm.requires_grad
    (lm * gout.to(cuda_device)).sum().backward()
    fast = (n_fft, hop) in ((1024, 256), (2048, 512))            # K6F / K6E: edge-strip zeroing + the backward kernel
    if os.environ.get("MIXFE_FORCE_GENERIC_BWD", "0") not in ("", "0"):
        fast = False
    if os.environ.get("MIXFE_FORCE_GENERIC", "0") not in ("", "0") and n_fft == 1024:
        fast = False                                             # K6F needs K1's tables
    assert pkg.launch_count() == n0 + (3 if fast else 2)         # forward kernel + K6F / K6
    for n in O.STEMS:
        got = leaves[n].grad[0].cpu().double()
        assert got.shape == ref64[n].shape
        scale = float(ref64[n].abs().max())
        err = float((got - ref64[n]).abs().max())
        err32 = float((ref32[n] - ref64[n]).abs().max())
        assert err <= max(4.0 * err32, 1e-4 * scale), f"{n}: err {err:.3g}, fp32 oracle err {err32:.3g}, scale {scale:.3g}"


def test_backward_unbatched_and_partial_requires_grad(cuda_device):
    """(2,T) stems as the per-clip callers pass them; only two stems need gradients."""
    cfg = O.FrontEndConfig()
    T = 12000
    clip = _noisy_clip(5, T)
    gout = torch.randn(1, 8, 128, cfg.n_frames(T), generator=torch.Generator().manual_seed(4))
    ref = _oracle_grad(clip, cfg, gout, torch.float64)
    pre = pkg.MelSpectrogramPreprocessor()
    leaves = {n: clip[n].to(cuda_device) for n in O.STEMS}
    leaves["bass"].requires_grad_(True)
    leaves["other"].requires_grad_(True)
    (pre(leaves) * gout.to(cuda_device)).sum().backward()
    assert leaves["vocals"].grad is None and leaves["drums"].grad is None
    for n in ("bass", "other"):
        got = leaves[n].grad.cpu().double()
        assert got.shape == (2, T)
        assert float((got - ref[n]).abs().max()) <= 1e-3 * float(ref[n].abs().max())


def test_no_grad_path_unchanged(cuda_device):
    pre = pkg.MelSpectrogramPreprocessor()
    clip = {n: v[None].to(cuda_device) for n, v in O.musical_clip(2, 8000).items()}
    with torch.no_grad():
        a = pre({n: v.clone().requires_grad_(True) for n, v in clip.items()})
    b = pre(clip)
    assert not a.requires_grad and torch.equal(a, b)


def test_batched_backward_equals_per_clip(cuda_device):
    pre = pkg.MelSpectrogramPreprocessor(44100, 2048, 512, 80)
    T = 16000
    clips = [_noisy_clip(40 + i, T) for i in range(3)]
    F = 1 + T // 512
    gout = torch.randn(3, 8, 80, F, generator=torch.Generator().manual_seed(9)).to(cuda_device)
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device).requires_grad_(True) for n in O.STEMS}
    (pre(batch) * gout).sum().backward()
    for i, c in enumerate(clips):
        one = {n: c[n][None].to(cuda_device).requires_grad_(True) for n in O.STEMS}
        (pre(one) * gout[i:i + 1]).sum().backward()
        for n in O.STEMS:
            d = float((batch[n].grad[i] - one[n].grad[0]).abs().max())
            assert d <= 1e-6 * float(one[n].grad.abs().max())     # atomics: order-dependent only at the clip edges


@pytest.mark.parametrize("T", [513, 700, 1025, 1536, 3327, 3328, 3329, 6655, 20000, 44100 + 17])
def test_fast_backward_lengths_and_generic_agreement(cuda_device, T):
    """K6F (1024/256, with the forward's log-mel) against the generic K6 on the same inputs, over clip lengths around
    its tiling (13 hop regions = 3328 samples per CTA) and the reflect-fold strips (513 samples at each end), and
    bitwise run-to-run determinism (every sample is written once; the strips take two commutative adds)."""
    clip = _noisy_clip(T, T)
    stems = [clip[n][None].to(cuda_device) for n in O.STEMS]
    plan = pkg.MelSpectrogramPreprocessor()._plan(cuda_device)
    _, lm = plan.features_logmel(stems, False, True)
    g = torch.randn(lm.shape, generator=torch.Generator().manual_seed(T)).to(cuda_device)
    slow = plan.logmel_backward(stems, g)
    fast = plan.logmel_backward(stems, g, logmel=lm)
    again = plan.logmel_backward(stems, g, logmel=lm)
    for a, b_, c in zip(slow, fast, again):
        scale = float(a.abs().max())
        assert float((a - b_).abs().max()) <= 2e-4 * scale, (T, float((a - b_).abs().max()), scale)
        assert torch.equal(b_, c)


def test_fast_backward_batched_and_strided(cuda_device):
    T = 9000
    clips = [_noisy_clip(70 + i, T) for i in range(3)]
    wide = torch.zeros(3, 8, T + 6, device=cuda_device)
    for i, c in enumerate(clips):
        for k, n in enumerate(O.STEMS):
            wide[i, 2 * k:2 * k + 2, 2:2 + T] = c[n].to(cuda_device)
    views = [wide[:, 2 * k:2 * k + 2, 2:2 + T] for k in range(4)]
    plan = pkg.MelSpectrogramPreprocessor()._plan(cuda_device)
    _, lm = plan.features_logmel(views, False, True)
    g = torch.randn(lm.shape, generator=torch.Generator().manual_seed(1)).to(cuda_device)
    fast = plan.logmel_backward(views, g, logmel=lm)
    for i in range(3):
        one = [v[i:i + 1].contiguous() for v in views]
        ref = plan.logmel_backward(one, g[i:i + 1], logmel=lm[i:i + 1])
        for a, b_ in zip(fast, ref):
            assert torch.equal(a[i:i + 1], b_)


@pytest.mark.parametrize("T", [1025, 1400, 2049, 3072, 6655, 6656, 6657, 13311, 20001, 44100 + 17])
def test_fast_backward_2048_lengths_and_generic_agreement(cuda_device, T):
    """K6E ((2048, 512), with the forward's log-mel) against the generic K6 on the same inputs, over clip lengths around its
    tiling (13 hop regions = 6656 samples per CTA), odd lengths (the kernel works in units of two samples) and the
    reflect-fold strips (1025 samples at each end); bitwise run-to-run determinism."""
    clip = _noisy_clip(T, T)
    stems = [clip[n][None].to(cuda_device) for n in O.STEMS]
    plan = pkg.MelSpectrogramPreprocessor(44100, 2048, 512, 80)._plan(cuda_device)
    _, lm = plan.features_logmel(stems, False, True)
    g = torch.randn(lm.shape, generator=torch.Generator().manual_seed(T)).to(cuda_device)
    n0 = pkg.launch_count()
    slow = plan.logmel_backward(stems, g)
    assert pkg.launch_count() - n0 == 1
    fast = plan.logmel_backward(stems, g, logmel=lm)
    assert pkg.launch_count() - n0 == 3                          # + strip zeroing + K6E
    again = plan.logmel_backward(stems, g, logmel=lm)
    for a, b_, c in zip(slow, fast, again):
        scale = float(a.abs().max())
        assert float((a - b_).abs().max()) <= 2e-4 * scale, (T, float((a - b_).abs().max()), scale)
        assert torch.equal(b_, c)


def test_fast_backward_2048_batched_strided_and_unaligned(cuda_device):
    """Rows that start on odd float offsets (no 8-byte alignment: the scalar load / store paths) inside a wider buffer."""
    T = 9001
    clips = [_noisy_clip(170 + i, T) for i in range(3)]
    wide = torch.zeros(3, 8, T + 7, device=cuda_device)
    for i, c in enumerate(clips):
        for k, n in enumerate(O.STEMS):
            wide[i, 2 * k:2 * k + 2, 3:3 + T] = c[n].to(cuda_device)
    views = [wide[:, 2 * k:2 * k + 2, 3:3 + T] for k in range(4)]
    plan = pkg.MelSpectrogramPreprocessor(44100, 2048, 512, 80)._plan(cuda_device)
    _, lm = plan.features_logmel(views, False, True)
    g = torch.randn(lm.shape, generator=torch.Generator().manual_seed(1)).to(cuda_device)
    fast = plan.logmel_backward(views, g, logmel=lm)
    slow = plan.logmel_backward(views, g)
    for a, b_ in zip(slow, fast):
        assert float((a - b_).abs().max()) <= 2e-4 * float(a.abs().max())
    for i in range(3):
        one = [v[i:i + 1].contiguous() for v in views]
        ref = plan.logmel_backward(one, g[i:i + 1], logmel=lm[i:i + 1])
        for a, b_ in zip(fast, ref):
            assert torch.equal(a[i:i + 1], b_)
