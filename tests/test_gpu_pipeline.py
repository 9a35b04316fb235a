"""GPU: HostPipeline (pinned host stems -> chunked H2D -> kernels -> D2H) returns exactly what the resident call
returns, for every chunk schedule shape (ramped ends, ragged tail, one chunk)."""
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from mixing_style_transfer_b200.pipeline import HostPipeline
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,chunk", [(13, 4), (7, 3), (5, 8), (1, 2), (24, 12)])
def test_pipeline_matches_resident(cuda_device, n, chunk):
    T = 9000
    fe = pkg.MixingFeatureExtractor()
    clips = [O.musical_clip(70 + i, T) for i in range(n)]
    host = {name: HostPipeline.pinned_like((n, 2, T)) for name in O.STEMS}
    for name in O.STEMS:
        host[name].copy_(torch.stack([c[name] for c in clips]))
    pipe = HostPipeline(fe, n, T, chunk=chunk, device=cuda_device)
    assert pipe.bounds[0][0] == 0 and pipe.bounds[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(pipe.bounds, pipe.bounds[1:]))
    assert max(hi - lo for lo, hi in pipe.bounds) <= min(chunk, n)
    hf = HostPipeline.pinned_like((n, 64)).fill_(-1)
    hl = HostPipeline.pinned_like((n, 8, 128, 1 + T // 256)).fill_(-1)
    for _ in range(2):                                   # second run reuses the slots
        pipe.run(host, hf, hl)
    f, lm = fe.features_and_logmel({k: v.to(cuda_device) for k, v in host.items()})
    assert torch.equal(hf, f.cpu()) and torch.equal(hl, lm.cpu())
    assert pipe.h2d_bytes == 4 * 8 * T * n and pipe.d2h_bytes == 4 * n * (64 + 8 * 128 * (1 + T // 256))
