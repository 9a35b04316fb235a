"""GPU parity of the augmentation kernels (K3 scan + K4/K5 FFT reverb) against the CPU oracle and the
golden waveforms recorded from the real reference.  Waveform tolerance:
|d| <= 1e-4 * max|ref|; the asserted bound (2e-5) is five times tighter and the achieved error is printed.
(K3 runs the 16 samples of a lane in fp32 and only the scanned states in fp64: the 500 Hz low-pass, poles at
r = 0.951, reaches 6e-6; everything else stays under 1e-6.)"""
import numpy as np
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O
from tests._util import golden

pytestmark = pytest.mark.gpu
G = golden()
TOL = 2e-5


def _decisions_from_oracle(par: O.ClipAug) -> pkg.ClipAugmentation:
    d = pkg.ClipAugmentation()
    for n in O.STEMS:
        s = par.stems[n]
        d.gain_db[n] = s.gain_db
        d.tilt[n] = s.tilt
        d.compress[n] = s.compress
        d.cutoff_hz[n] = s.cutoff_hz
    d.impulse = par.impulse
    return d


def _rel_err(got, ref):
    scale = max(float(ref[n].abs().max()) for n in O.STEMS) + 1e-30
    return max(float((got[n].cpu() - ref[n]).abs().max()) for n in O.STEMS) / scale


def _run(clip, par, dev):
    aug = pkg.AudioAugmenter()
    return aug.apply({n: v.to(dev) for n, v in clip.items()}, [_decisions_from_oracle(par)])


def _single(**kw):
    par = O.ClipAug()
    for n in O.STEMS:
        par.stems[n] = O.StemAug()
    for k, v in kw.items():
        setattr(par.stems["vocals"], k, v)
        setattr(par.stems["drums"], k, v)
    return par


@pytest.mark.parametrize("T", [30000, 70001])
@pytest.mark.parametrize("effect", ["none", "gain", "tilt_high", "tilt_low", "compress", "bw4000", "bw11999", "chain"])
def test_single_effects(cuda_device, effect, T):
    clip = O.musical_clip(17, T)
    par = {
        "none": _single(),
        "gain": _single(gain_db=torch.tensor([-6.5])),
        "tilt_high": _single(tilt="high"),
        "tilt_low": _single(tilt="low"),
        "compress": _single(compress=True),
        "bw4000": _single(cutoff_hz=4000.0),
        "bw11999": _single(cutoff_hz=11999.0),
        "chain": _single(gain_db=torch.tensor([7.25]), tilt="low", compress=True, cutoff_hz=6543.21),
    }[effect]
    got = _run(clip, par, cuda_device)
    ref = O.apply_clip_aug(clip, par)
    err = _rel_err(got, ref)
    print(f"{effect} T={T}: rel max err {err:.3g}")
    assert err < TOL, err
    for n in O.STEMS:
        assert got[n].shape == clip[n].shape and got[n].dtype == torch.float32
    if effect == "none":
        for n in O.STEMS:
            assert torch.equal(got[n].cpu(), clip[n])             # clone semantics (:386), bit exact


@pytest.mark.parametrize("seed,prob", [(1234, 1.0), (99, 0.5), (7, 0.5), (11, 0.5)])
def test_full_chain_vs_reference_golden(cuda_device, seed, prob):
    """Same seed -> same host decisions as the reference -> compare with the reference's own output."""
    T = 30000
    clip = O.musical_clip(seed, T)
    aug = pkg.AudioAugmenter(prob=prob)
    torch.manual_seed(seed)
    got = aug.augment_stems({n: v.to(cuda_device) for n, v in clip.items()}) if False else None
    # the reference draws the impulse noise on the audio's device; the golden was made on CPU, so draw on CPU
    torch.manual_seed(seed)
    dec = aug.draw(impulse_device="cpu")
    got = aug.apply({n: v.to(cuda_device) for n, v in clip.items()}, [dec])
    scale = max(float(np.abs(G[f"augment/{seed}/{prob}/{T}/{n}/sub5"]).max()) for n in O.STEMS)
    worst = 0.0
    for n in O.STEMS:
        ref = torch.from_numpy(G[f"augment/{seed}/{prob}/{T}/{n}/sub5"])
        worst = max(worst, float((got[n][:, ::5].cpu() - ref).abs().max()) / scale)
    print(f"seed {seed} prob {prob}: rel max err vs reference {worst:.3g}")
    assert worst < 2e-5, worst          # the reference's 22 050-tap fp32 conv1d carries ~5e-6 of its own noise


def test_reverb_only_and_energy_shares(cuda_device):
    T = 50000
    clip = O.musical_clip(23, T)
    par = _single()
    g = torch.Generator().manual_seed(5)
    t = torch.linspace(0, 0.5, 22050)
    par.impulse = torch.exp(-t / 0.125) * torch.randn(22050, generator=g) * 0.1
    got = _run(clip, par, cuda_device)
    ref = O.apply_clip_aug(clip, par, fft_reverb=True)
    err = _rel_err(got, ref)
    print(f"reverb only: rel max err {err:.3g}")
    assert err < TOL, err


def test_full_length_clip_chunked_scan_is_exact_enough(cuda_device):
    """10 s clip: 54 chunks per row, each warm-started; the LP-500 Hz tilt has the slowest poles (r = 0.951)."""
    T = 441000
    clip = O.synthetic_clip(31, T)
    torch.manual_seed(1234)
    par = O.draw_clip_aug(prob=1.0)
    got = _run(clip, par, cuda_device)
    ref = O.apply_clip_aug(clip, par, fft_reverb=True)
    err = _rel_err(got, ref)
    print(f"10 s clip, all effects: rel max err {err:.3g}")
    assert err < TOL, err


def test_long_clip_two_reverb_segments(cuda_device):
    T = 1_100_000                                   # > 2^20 - 22050 + 1: overlap-save with two segments
    clip = {n: v * 0.5 for n, v in O.synthetic_clip(41, T).items()}
    par = _single(tilt="high")
    g = torch.Generator().manual_seed(6)
    t = torch.linspace(0, 0.5, 22050)
    par.impulse = torch.exp(-t / 0.125) * torch.randn(22050, generator=g) * 0.1
    got = _run(clip, par, cuda_device)
    ref = O.apply_clip_aug(clip, par, fft_reverb=True)
    err = _rel_err(got, ref)
    print(f"two segments: rel max err {err:.3g}")
    assert err < TOL, err


def test_batched_clips_with_individual_decisions(cuda_device):
    T = 20000
    clips = [O.musical_clip(s, T) for s in (1, 2, 3, 4)]
    pars = []
    for s in (1, 2, 3, 4):
        torch.manual_seed(100 + s)
        pars.append(O.draw_clip_aug(prob=0.5))
    batch = {n: torch.stack([c[n] for c in clips]).to(cuda_device) for n in O.STEMS}
    aug = pkg.AudioAugmenter()
    n0 = pkg.launch_count()
    got = aug.apply(batch, [_decisions_from_oracle(p) for p in pars])
    got2 = aug.apply(batch, [_decisions_from_oracle(p) for p in pars])
    assert pkg.launch_count() - n0 in (4, 14)                     # parameter fetch + K3 (+ 5 reverb launches) per call
    for n in O.STEMS:
        assert got[n].shape == (4, 2, T) and torch.equal(got[n], got2[n])
    for i, (c, p) in enumerate(zip(clips, pars)):
        ref = O.apply_clip_aug(c, p, fft_reverb=True)
        err = _rel_err({n: got[n][i] for n in O.STEMS}, ref)
        assert err < TOL, (i, err)


def test_apply_writes_into_a_slice_of_the_batch(cuda_device):
    """`out=`: the negatives are written straight into their slice of the tensor the anchors live in."""
    T = 20000
    clips = [O.musical_clip(s, T) for s in (5, 6)]
    pars = []
    for s in (5, 6):
        torch.manual_seed(200 + s)
        pars.append(O.draw_clip_aug(prob=1.0))
    batch = {n: torch.zeros(6, 2, T, device=cuda_device) for n in O.STEMS}
    for n in O.STEMS:
        batch[n][0:2] = torch.stack([c[n] for c in clips]).to(cuda_device)
    aug = pkg.AudioAugmenter()
    dec = [_decisions_from_oracle(p) for p in pars]
    ret = aug.apply({n: batch[n][0:2] for n in O.STEMS}, dec, out={n: batch[n][4:6] for n in O.STEMS})
    fresh = aug.apply({n: batch[n][0:2] for n in O.STEMS}, dec)
    for n in O.STEMS:
        assert ret[n].data_ptr() == batch[n][4:6].data_ptr()
        assert torch.equal(batch[n][4:6], fresh[n])
        assert float(batch[n][2:4].abs().max()) == 0.0                    # neighbours untouched
    with pytest.raises(ValueError):
        aug.apply({n: batch[n][0:2] for n in O.STEMS}, dec, out={n: torch.empty(2, 2, T + 1, device=cuda_device) for n in O.STEMS})


def test_augment_stems_draws_like_the_reference(cuda_device):
    """augment_stems consumes the global RNG in the reference's order (checked on CPU in test_host_logic);
    here: same seed twice -> identical output, inputs untouched."""
    clip = {n: v.to(cuda_device) for n, v in O.musical_clip(8, 16000).items()}
    keep = {n: v.clone() for n, v in clip.items()}
    aug = pkg.AudioAugmenter(prob=0.7)
    torch.manual_seed(3)
    a = aug.augment_stems(clip)
    torch.manual_seed(3)
    b = aug.augment_stems(clip)
    for n in O.STEMS:
        assert torch.equal(a[n], b[n]) and torch.equal(clip[n], keep[n])
    assert list(a.keys()) == list(clip.keys())


def test_bad_arguments(cuda_device):
    aug = pkg.AudioAugmenter()
    clip = {n: v.to(cuda_device) for n, v in O.synthetic_clip(0, 4096).items()}
    with pytest.raises(ValueError):
        aug.apply(clip, [])
    d = pkg.ClipAugmentation()
    for n in O.STEMS:
        d.gain_db[n] = None; d.tilt[n] = None; d.compress[n] = False; d.cutoff_hz[n] = None
    arr = aug._pack([d])
    arr[0].tilt_on[0] = 1
    for j, v in enumerate([1.0, 0.0, 0.0, 1.0, -2.1, 1.1]):           # poles outside the unit circle
        arr[0].tilt_sos[0][j] = v
    st = [clip[n][None] for n in O.STEMS]
    with pytest.raises(ValueError, match="unstable"):
        aug._plan(cuda_device).augment(st, arr, None)


def _fuzz_lengths():
    import random
    rnd = random.Random(77)
    edge = [1, 2, 3, 31, 32, 33, 127, 511, 512, 513, 1023, 1025, 8191, 8193]          # around the scan's block sizes
    return [(i, T) for i, T in enumerate(edge + [rnd.randrange(600, 90000) for _ in range(8)])]


@pytest.mark.parametrize("seed,T", _fuzz_lengths())
def test_random_decisions_lengths_and_unaligned_views(cuda_device, seed, T):
    """Seeded random augmentation decisions (prob 0.7) on clips of awkward lengths, the stems passed as views that
    start 1..3 floats into a wider tensor (no 128-bit loads possible)."""
    torch.manual_seed(1000 + seed)
    par = O.draw_clip_aug(prob=0.7)
    if T < 22050:
        par.impulse = None                     # F.conv1d 'same' against a 22050-tap impulse needs nothing special,
                                               # but the oracle's direct form is slow; the reverb has its own tests
    clip = O.musical_clip(300 + seed, T)
    off = 1 + seed % 3
    wide = {n: torch.zeros(2, T + 8).to(cuda_device) for n in O.STEMS}
    views = {}
    for n in O.STEMS:
        wide[n][:, off:off + T] = clip[n].to(cuda_device)
        views[n] = wide[n][:, off:off + T]
    aug = pkg.AudioAugmenter()
    got = aug.apply(views, [_decisions_from_oracle(par)])
    ref = O.apply_clip_aug(clip, par, fft_reverb=True)
    err = _rel_err(got, ref)
    assert err < TOL, f"T={T} off={off}: rel max err {err:.3g}"
    for n in O.STEMS:                                              # inputs untouched, padding untouched
        assert torch.equal(wide[n][:, off:off + T].cpu(), clip[n])
        assert float(wide[n][:, :off].abs().max()) == 0.0 and float(wide[n][:, off + T:].abs().max()) == 0.0
