"""Plans of different configurations living side by side on one device: the dynamic
shared-memory opt-in is state of (kernel, device), so a later, smaller plan must not lower the limit an
earlier plan launches with.  Sequence taken from INTEGRATION.md: encoder forward, MR-STFT loss forward /
backward (which builds (1024,256,64)-style and n_fft=512/2048 plans lazily), encoder again."""
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu


def test_smaller_plans_do_not_break_earlier_ones(cuda_device):
    T = 30000
    clip = {n: v.to(cuda_device) for n, v in O.musical_clip(3, T).items()}
    batch = {n: v[None] for n, v in clip.items()}
    fe = pkg.MixingFeatureExtractor()                             # K1, 128 mels: the largest shared-memory map
    enc = pkg.MelSpectrogramPreprocessor()
    f0 = fe.extract_all_features(clip)
    lm0 = enc(batch)
    # smaller footprints of the same kernels: fewer mels (K1), other n_fft (K1G / K6 / K8)
    small = pkg.MixingFeatureExtractor(44100, 1024, 256, 64)
    small.extract_all_features(clip)
    ev = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)
    ev.extract_all_features(clip)
    tiny = pkg.MixingFeatureExtractor(44100, 256, 64, 32)
    tiny.extract_all_features(clip)
    loss = pkg.MultiResolutionSTFTLoss()
    x = torch.stack([clip["vocals"], clip["bass"]]).requires_grad_(True)
    y = torch.stack([clip["drums"], clip["other"]])
    loss(x, y).backward()
    # the first plans again: forward, and the backward that opts K6 in
    f1 = fe.extract_all_features(clip)
    lm1 = enc(batch)
    assert torch.equal(f0, f1) and torch.equal(lm0, lm1)
    req = {n: v.clone().requires_grad_(True) for n, v in batch.items()}
    enc(req).sum().backward()
    g_big = req["vocals"].grad.clone()
    enc_small = pkg.MelSpectrogramPreprocessor(44100, 512, 128, 40)
    req2 = {n: v.clone().requires_grad_(True) for n, v in batch.items()}
    enc_small(req2).sum().backward()
    req3 = {n: v.clone().requires_grad_(True) for n, v in batch.items()}
    enc(req3).sum().backward()
    assert torch.equal(g_big, req3["vocals"].grad)
    torch.cuda.synchronize()


def test_host_pipeline_default_device(cuda_device):
    from mixing_style_transfer_b200.pipeline import HostPipeline
    T, n = 8192, 3
    fe = pkg.MixingFeatureExtractor()
    pipe = HostPipeline(fe, n, T, chunk=2)                        # device=None: the current CUDA device
    assert pipe.device == torch.device("cuda", torch.cuda.current_device())
    assert pipe.plan is not fe._plan(pipe.device)                 # private plan: the pipeline runs on its own stream
    host = {s: HostPipeline.pinned_like((n, 2, T)) for s in pkg.STEMS}
    clips = [O.musical_clip(40 + i, T) for i in range(n)]
    for s in pkg.STEMS:
        host[s].copy_(torch.stack([c[s] for c in clips]))
    hf = HostPipeline.pinned_like((n, 64))
    hl = HostPipeline.pinned_like((n, 8, 128, 1 + T // 256))
    pipe.run(host, hf, hl)
    ref_f, ref_lm = fe.features_and_logmel({s: host[s].to(cuda_device) for s in pkg.STEMS})
    assert torch.equal(hf, ref_f.cpu()) and torch.equal(hl, ref_lm.cpu())


def test_output_buffers_are_validated(cuda_device):
    fe = pkg.MixingFeatureExtractor()
    plan = fe._plan(cuda_device)
    T = 4096
    stems = [torch.zeros(2, 2, T, device=cuda_device) for _ in range(4)]
    F = 1 + T // 256
    with pytest.raises(ValueError, match="features_out"):
        plan.features_logmel(stems, True, False, features_out=torch.empty(1, 64, device=cuda_device))
    with pytest.raises(ValueError, match="features_out"):
        plan.features_logmel(stems, True, False, features_out=torch.empty(2, 64, device=cuda_device, dtype=torch.float64))
    with pytest.raises(ValueError, match="features_out"):
        plan.features_logmel(stems, True, False, features_out=torch.empty(2, 128, device=cuda_device)[:, ::2])
    with pytest.raises(ValueError, match="logmel_out"):
        plan.features_logmel(stems, False, True, logmel_out=torch.empty(2, 8, 128, F + 1, device=cuda_device))
    with pytest.raises(ValueError, match="logmel_out"):
        plan.features_logmel(stems, False, True, logmel_out=torch.empty(2, 8, 128, F))
    aug = pkg.AudioAugmenter(prob=0.0)
    buf = {n: torch.zeros(3, 2, T, device=cuda_device) for n in pkg.STEMS}
    d = [aug.draw(), aug.draw()]
    with pytest.raises(ValueError, match="overlaps"):
        aug.apply({n: v[0:2] for n, v in buf.items()}, d, out={n: v[1:3] for n, v in buf.items()})


def test_work_counters_survive_changing_grids(cuda_device):
    """K1 / K1E hand out tiles through two per-plan counters that the last CTA of a launch rezeroes: launches whose grids
    differ (1 clip, many clips, SMs reserved, short clips with fewer tiles than SMs) must leave them clean for the next."""
    T = 30000
    clips = [O.musical_clip(800 + i, T) for i in range(5)]
    for cfg in ((44100, 1024, 256, 128), (44100, 2048, 512, 80)):
        fe = pkg.MixingFeatureExtractor(*cfg)
        plan = fe._plan(cuda_device)
        batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
        f_ref, lm_ref = fe.features_and_logmel(batch)
        for rounds in range(3):
            for k in (1, 5, 2):
                sub = {n: v[:k] for n, v in batch.items()}
                f, lm = fe.features_and_logmel(sub)
                assert torch.equal(f, f_ref[:k]) and torch.equal(lm, lm_ref[:k])
            plan.reserve_sms(100)                        # a much smaller persistent grid
            f, lm = fe.features_and_logmel(batch)
            plan.reserve_sms(0)
            assert torch.equal(f, f_ref) and torch.equal(lm, lm_ref)
            short = {n: v[:1, :, :3000] for n, v in batch.items()}      # fewer tiles than SMs
            f1, _ = fe.features_and_logmel(short)
            f2, _ = fe.features_and_logmel(short)
            assert torch.equal(f1, f2)
