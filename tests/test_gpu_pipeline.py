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


def test_pipeline_one_allocation_matches_resident(cuda_device):
    """The four stems as equally spaced views of one pinned buffer: one strided transfer per chunk."""
    T, n = 9000, 7
    fe = pkg.MixingFeatureExtractor()
    clips = [O.musical_clip(170 + i, T) for i in range(n)]
    whole = HostPipeline.pinned_like((4, n, 2, T))
    host = {name: whole[k] for k, name in enumerate(O.STEMS)}
    for name in O.STEMS:
        host[name].copy_(torch.stack([c[name] for c in clips]))
    pipe = HostPipeline(fe, n, T, chunk=3, device=cuda_device)
    hf = HostPipeline.pinned_like((n, 64)).fill_(-1)
    hl = HostPipeline.pinned_like((n, 8, 128, 1 + T // 256)).fill_(-1)
    pipe.run(host, hf, hl)
    f, lm = fe.features_and_logmel({k: v.to(cuda_device) for k, v in host.items()})
    assert torch.equal(hf, f.cpu()) and torch.equal(hl, lm.cpu())


def test_prefetching_submit_wait_matches_run(cuda_device):
    """Three different batches queued with one batch of prefetch (submit k+1 before wait k, two sets of host outputs,
    odd number of chunks per batch so that the slot parity flips between batches) give what the blocking run gives."""
    T, n = 9000, 7
    fe = pkg.MixingFeatureExtractor()
    pipe = HostPipeline(fe, n, T, chunk=3, device=cuda_device)
    assert len(pipe.bounds) % 2 == 1
    hosts = []
    for k in range(3):
        clips = [O.musical_clip(200 + 10 * k + i, T) for i in range(n)]
        h = {name: HostPipeline.pinned_like((n, 2, T)) for name in O.STEMS}
        for name in O.STEMS:
            h[name].copy_(torch.stack([c[name] for c in clips]))
        hosts.append(h)
    want = []
    for h in hosts:
        hf = HostPipeline.pinned_like((n, 64)); hl = HostPipeline.pinned_like((n, 8, 128, 1 + T // 256))
        pipe.run(h, hf, hl)
        want.append((hf.clone(), hl.clone()))
    outs = [(HostPipeline.pinned_like((n, 64)).fill_(-1), HostPipeline.pinned_like((n, 8, 128, 1 + T // 256)).fill_(-1))
            for _ in range(2)]
    got = []
    pending = pipe.submit(hosts[0], *outs[0])
    for k in range(1, 3):
        nxt = pipe.submit(hosts[k], *outs[k & 1])
        pipe.wait(pending)
        got.append(tuple(t.clone() for t in outs[(k - 1) & 1]))
        pending = nxt
    pipe.wait(pending)
    got.append(tuple(t.clone() for t in outs[0]))
    for (wf, wl), (gf, gl) in zip(want, got):
        assert torch.equal(wf, gf) and torch.equal(wl, gl)


@pytest.mark.parametrize("one_alloc", [False, True])
@pytest.mark.parametrize("G,chunk", [(5, 2), (9, 2), (3, 4), (11, 4)])
def test_triplet_pipeline_matches_resident(cuda_device, G, chunk, one_alloc):
    """TripletPipeline (host anchors + positives -> device negatives -> features + log-mel of 3G clips) against the
    resident calls on the same decisions; blocking run and prefetching submit/wait."""
    from mixing_style_transfer_b200.pipeline import TripletPipeline
    T = 30000
    fe = pkg.MixingFeatureExtractor()
    aug = pkg.AudioAugmenter(prob=0.6)
    clips = [O.musical_clip(300 + i, T) for i in range(2 * G)]
    if one_alloc:        # the four stems as equally spaced views of one pinned buffer: one strided transfer per chunk
        whole = HostPipeline.pinned_like((4, 2 * G, 2, T))
        host = {name: whole[k] for k, name in enumerate(O.STEMS)}
    else:                # separate tensors: one strided transfer per stem
        host = {name: HostPipeline.pinned_like((2 * G, 2, T)) for name in O.STEMS}
    for name in O.STEMS:
        host[name].copy_(torch.stack([c[name] for c in clips]))
    torch.manual_seed(11)
    decs = [aug.draw_batch(G, impulse_device=cuda_device) for _ in range(2)]
    pipe = TripletPipeline(fe, aug, G, T, chunk=chunk, device=cuda_device)
    assert pipe.bounds[0][0] == 0 and pipe.bounds[-1][1] == G
    F = 1 + T // 256
    want = []
    for dec in decs:
        dev = {k: v.to(cuda_device) for k, v in host.items()}
        neg = aug.apply({k: v[:G] for k, v in dev.items()}, dec)
        full = {k: torch.cat([dev[k], neg[k]]) for k in O.STEMS}
        f, lm = fe.features_and_logmel(full)
        want.append((f.cpu(), lm.cpu()))
    outs = [(HostPipeline.pinned_like((3 * G, 64)).fill_(-1), HostPipeline.pinned_like((3 * G, 8, 128, F)).fill_(-1))
            for _ in range(2)]
    pipe.run(host, decs[0], *outs[0])
    assert torch.equal(outs[0][0], want[0][0]) and torch.equal(outs[0][1], want[0][1])
    outs[0][0].fill_(-1); outs[0][1].fill_(-1)
    t0 = pipe.submit(host, decs[0], *outs[0])
    t1 = pipe.submit(host, decs[1], *outs[1])
    pipe.wait(t0)
    assert torch.equal(outs[0][0], want[0][0]) and torch.equal(outs[0][1], want[0][1])
    pipe.wait(t1)
    assert torch.equal(outs[1][0], want[1][0]) and torch.equal(outs[1][1], want[1][1])
    with pytest.raises(ValueError):
        pipe.run(host, decs[0].slice(0, G - 1) if G > 1 else [], *outs[0])


def test_triplet_pipeline_outputs_may_stay_on_the_device(cuda_device):
    """feat_out / lm_out may be CUDA tensors (the encoder consumes the log-mels on the GPU: only the stems cross PCIe)."""
    from mixing_style_transfer_b200.pipeline import TripletPipeline
    T, G = 20000, 5
    fe = pkg.MixingFeatureExtractor()
    aug = pkg.AudioAugmenter(prob=0.6)
    clips = [O.musical_clip(400 + i, T) for i in range(2 * G)]
    whole = HostPipeline.pinned_like((4, 2 * G, 2, T))
    host = {name: whole[k] for k, name in enumerate(O.STEMS)}
    for name in O.STEMS:
        host[name].copy_(torch.stack([c[name] for c in clips]))
    torch.manual_seed(5)
    dec = aug.draw_batch(G, impulse_device=cuda_device)
    pipe = TripletPipeline(fe, aug, G, T, chunk=2, device=cuda_device)
    F = 1 + T // 256
    hf, hl = HostPipeline.pinned_like((3 * G, 64)), HostPipeline.pinned_like((3 * G, 8, 128, F))
    pipe.run(host, dec, hf, hl)
    df = torch.full((3 * G, 64), -1.0, device=cuda_device)
    dl = torch.full((3 * G, 8, 128, F), -1.0, device=cuda_device)
    pipe.wait(pipe.submit(host, dec, df, dl))
    assert torch.equal(df.cpu(), hf) and torch.equal(dl.cpu(), hl)
