"""CPU: the BS.1770-4 oracle (oracle/bs1770.py) against the only published vectors this extension has -- the 48 kHz
K-weighting table of ITU-R BS.1770-4 Annex 1 -- and against known answers of the definition."""
import numpy as np
import pytest

from oracle import bs1770 as B


def test_k_weighting_reproduces_the_itu_48k_table():
    sos = B.k_weighting_sos(48000)
    np.testing.assert_allclose(sos[0, :3], B.ITU_48K_SHELF[0], rtol=0, atol=1e-9)
    np.testing.assert_allclose(sos[0, 3:], B.ITU_48K_SHELF[1], rtol=0, atol=1e-9)
    np.testing.assert_allclose(sos[1, :3], B.ITU_48K_HIGHPASS[0], rtol=0, atol=1e-12)
    np.testing.assert_allclose(sos[1, 3:], B.ITU_48K_HIGHPASS[1], rtol=0, atol=1e-9)


def test_full_scale_997hz_sine_reads_minus_3_lufs_mono_and_0_stereo():
    """BS.1770-4: a 0 dBFS 997 Hz sine in one channel reads -3.01 LKFS."""
    for fs in (48000, 44100):
        t = np.arange(5 * fs) / fs
        s = np.sin(2 * np.pi * 997 * t)
        assert abs(B.integrated_loudness(s[None], fs) + 3.01) < 0.01
        assert abs(B.integrated_loudness(np.stack([s, s]), fs)) < 0.01


def test_gates():
    fs = 44100
    t = np.arange(10 * fs) / fs
    loud = 0.1 * np.sin(2 * np.pi * 997 * t)
    x = loud.copy()
    x[5 * fs:] *= 1e-3                                   # 60 dB down: removed by the relative gate
    ref = B.integrated_loudness(loud[None, :5 * fs], fs)
    got = B.integrated_loudness(x[None], fs)
    # without the gate the quiet half would pull the mean down by 3 dB; with it only the three blocks that
    # straddle the edge (>= 25 % loud, i.e. within 10 LU) remain in the average
    assert ref - 0.2 < got < ref
    assert B.integrated_loudness(np.zeros((2, fs)), fs) == float("-inf")   # nothing passes the absolute gate
    with pytest.raises(ValueError):
        B.integrated_loudness(np.zeros((2, fs // 10 * 4 - 1)), fs)
    with pytest.raises(ValueError):
        B.integrated_loudness(np.zeros((2, 44101)), 44101)


def test_short_term_iqr_of_a_steady_tone_is_zero_and_of_a_level_step_is_the_step():
    """DR_IQR = IQR(L_short): a steady sine has one short-term level; a sine that drops by 12 dB half way through has
    two equally populated levels (plus the few windows that straddle the step), so the IQR is close to 12 LU."""
    fs, T = 48000, 48000 * 20
    t = np.arange(T) / fs
    tone = 0.3 * np.sin(2 * np.pi * 997.0 * t)
    steady = np.stack([tone, tone])
    assert abs(B.short_term_iqr(steady, fs)) < 1e-6
    env = np.where(t < 10.0, 1.0, 10.0 ** (-12.0 / 20.0))
    stepped = np.stack([tone * env, tone * env])
    iqr = B.short_term_iqr(stepped, fs)
    assert 11.5 < iqr <= 12.0 + 1e-6, iqr
    l = B.short_term_loudness(stepped, fs)
    assert l.shape == ((T - 3 * fs) // (fs // 10) + 1,)
    assert abs((l[0] - l[-1]) - 12.0) < 1e-3


def test_short_term_iqr_drops_silent_windows_and_needs_three_seconds():
    fs = 44100
    x = np.zeros((2, fs * 8))
    x[:, :fs * 4] = 0.2 * np.sin(2 * np.pi * 440.0 * np.arange(fs * 4) / fs)
    l = B.short_term_loudness(x, fs)
    assert (l < -70.0).any()                                  # the silent tail
    assert B.short_term_iqr(x, fs) < 15.0                     # bounded by the audible windows, not by -inf
    with pytest.raises(ValueError, match="3 s"):
        B.short_term_loudness(np.zeros((2, fs * 3 - 1)), fs)
