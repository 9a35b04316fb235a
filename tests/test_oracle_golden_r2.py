"""CPU: the oracle's single-effect restatements against round-2 golden vectors of the real reference
(tests/golden/make_golden_r2.py), and the song_A known-answer vector (BASELINE config 1) replayed from the
committed copy of the reference's asset."""
import os

import numpy as np
import pytest
import torch

from oracle import mixfe_oracle as O
from tests._util import golden

HERE = os.path.dirname(os.path.abspath(__file__))
G2 = np.load(os.path.join(HERE, "golden", "reference_outputs_r2.npz"))
X = O.musical_clip(77, 30000)["drums"]


def _rel(got, ref):
    ref = torch.from_numpy(ref)
    return float((got - ref).abs().max()) / float(ref.abs().max())


@pytest.mark.parametrize("seed,decay", [(5, 0.5), (6, 0.25), (7, 0.1)])
def test_apply_reverb_draw_and_value(seed, decay):
    torch.manual_seed(seed)
    h = O.draw_impulse(decay)
    assert h.shape == (int(44100 * decay),)
    for fft in (True, False):
        err = _rel(O.reverb_standalone(X, h, fft=fft), G2[f"apply_reverb/{seed}/{decay}"])
        assert err < 2e-5, (fft, err)          # the reference's 22 050-tap fp32 conv1d carries ~1e-5 of its own rounding


@pytest.mark.parametrize("thr,ratio", [(-20, 4), (-30, 2.5), (-12, 8), (-6.5, 3)])
def test_apply_compression_arguments(thr, ratio):
    got = O.compress_static(X, thr, ratio)
    ref = torch.from_numpy(G2[f"apply_compression/{thr}/{ratio}"])
    assert float((got - ref).abs().max()) <= 2e-7 * float(ref.abs().max())    # same fp32 formula; torch's vector / scalar log10 paths


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_tilt_and_bandwidth_draws(seed):
    torch.manual_seed(seed)
    kind = "high" if torch.rand(1) < 0.5 else "low"
    assert _rel(O.spectral_tilt(X, kind), G2[f"apply_spectral_tilt/{seed}"]) < 1e-6
    torch.manual_seed(seed)
    fc = (torch.rand(1) * 8000 + 4000).item()
    assert _rel(O.bandwidth_limit(X, fc), G2[f"apply_bandwidth_limit/{seed}"]) < 1e-6


def test_both_tilt_branches_are_covered():
    kinds = set()
    for seed in (1, 2, 3, 4):
        torch.manual_seed(seed)
        kinds.add(bool(torch.rand(1) < 0.5))
    assert kinds == {True, False}


def test_song_A_known_answer_vector():
    """BASELINE config 1: the reference's own asset as a synthetic 4-stem 10 s clip; features and log-mel statistics
    recorded from the real reference (tests/golden/make_golden.py)."""
    G = golden()
    pcm = np.load(os.path.join(HERE, "golden", "song_A_int16.npz"))["pcm"]
    assert pcm.shape == (661538, 2) and pcm.dtype == np.int16
    clip = O.song_a_clip(pcm)
    got = O.mixing_features(clip)
    ref = torch.from_numpy(G["features/song_A"])
    assert float(((got - ref).abs() - (1e-5 * ref.abs() + 2e-6)).max()) <= 0.0
    # SURVEY section 8c prints the same vector to 7 digits
    assert abs(float(got[0]) - 0.04255972) < 1e-7 and abs(float(got[30]) - 0.6019827) < 1e-6
    lm = O.log_mel({n: v[None] for n, v in clip.items()})
    st = G["logmel/song_A/stats"]
    assert lm.shape == (1, 8, 128, 1723)
    assert abs(float(lm.double().mean()) - st[0]) < 1e-5 and float(lm.min()) == pytest.approx(st[1], abs=1e-6)
    assert float(lm.max()) == pytest.approx(st[2], abs=1e-5)
    assert float(lm[0, 0, 64, 861]) == pytest.approx(st[3], abs=2e-5)
    assert float(lm[0, 7, 127, 1722]) == pytest.approx(st[4], abs=2e-4)
