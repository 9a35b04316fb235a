"""Run-to-run determinism on the GPU: every kernel that promises fixed-order reductions must give bit-identical
results across repeated launches (a missing barrier or a shared-memory race shows up here as a flaky bit).
compute-sanitizer is not available on the GPU pool, so this is the race check the suite has."""
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu


def _batch(dev, n, T, seed=0):
    clips = [O.musical_clip(seed + i, T) for i in range(n)]
    return {name: torch.stack([c[name] for c in clips]).to(dev) for name in O.STEMS}


@pytest.mark.parametrize("cfg", [(1024, 256, 128), (2048, 512, 80), (512, 128, 64)])
def test_features_logmel_bitwise_repeatable(cuda_device, cfg):
    fe = pkg.MixingFeatureExtractor(44100, *cfg)
    batch = _batch(cuda_device, 24, 50000 + cfg[0] + 3)          # odd T: scalar staging on the right channel
    f0, l0 = fe.features_and_logmel(batch)
    f0, l0 = f0.clone(), l0.clone()
    for _ in range(15):
        f, l = fe.features_and_logmel(batch)
        assert torch.equal(f, f0) and torch.equal(l, l0)


def test_features_logmel_bitwise_repeatable_aligned_long(cuda_device):
    """The TMA paths (staging + statistics rows) on a batch big enough to fill the persistent grid several times."""
    fe = pkg.MixingFeatureExtractor()
    g = torch.Generator(device=cuda_device).manual_seed(5)
    batch = {n: (torch.rand(40, 2, 220500, device=cuda_device, generator=g) - 0.5) * 0.4 for n in O.STEMS}
    f0, l0 = fe.features_and_logmel(batch)
    f0, l0 = f0.clone(), l0.clone()
    for _ in range(10):
        f, l = fe.features_and_logmel(batch)
        assert torch.equal(f, f0) and torch.equal(l, l0)
    # and the same clips one at a time (other tile -> CTA assignment) give the same bits
    for i in (0, 17, 39):
        f1, l1 = fe.features_and_logmel({n: v[i:i + 1] for n, v in batch.items()})
        assert torch.equal(f1[0], f0[i]) and torch.equal(l1[0], l0[i])


def test_augment_bitwise_repeatable(cuda_device):
    aug = pkg.AudioAugmenter(prob=1.0)
    batch = _batch(cuda_device, 6, 60000, seed=50)
    torch.manual_seed(11)
    decisions = [aug.draw(impulse_device=cuda_device) for _ in range(6)]
    ref = {n: v.clone() for n, v in aug.apply(batch, decisions).items()}
    for _ in range(8):
        out = aug.apply(batch, decisions)
        for n in O.STEMS:
            assert torch.equal(out[n], ref[n])
