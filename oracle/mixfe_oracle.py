"""CPU restatement of the reference DSP front end (TEST INFRASTRUCTURE ONLY).

What is restated, and where it lives in the reference (paths relative to the
reference checkout; torchaudio paths point into torchaudio 2.11.0, scipy paths
into scipy 1.18.1 -- third-party code the reference calls but does not vendor):

* log-mel front end          src/model.py:17-67  (MelSpectrogramPreprocessor)
* 64-D mixing features       src/mixing_utils.py:16-357 (MixingFeatureExtractor)
* negative-pair augmentation src/mixing_utils.py:364-479 (AudioAugmenter)
* mel spectrogram            torchaudio transforms/_transforms.py:566-620,
                             functional/functional.py:54-145 (spectrogram),
                             :518-587 (melscale_fbanks)
* Butterworth design, DF2T   scipy.signal.butter / sosfilt (called at
                             src/mixing_utils.py:426-432, 452-455)

The STFT itself is delegated to ``torch.stft`` -- the very kernel the reference
reaches through torchaudio -- so the fp32 oracle is bit-identical to the
reference on the same host.  ``dtype=torch.float64`` gives the "exact" variant
used to measure how far the fp32 reference itself is from the true value.

Parity: pinned (see oracle/__init__.py).  Nothing in the product imports this.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

STEMS = ("vocals", "bass", "drums", "other")       # src/model.py:54, src/mixing_utils.py:287
_SORTED_STEMS = tuple(sorted(STEMS))               # _flatten_features sorts keys, :324


@dataclass(frozen=True)
class FrontEndConfig:
    """Constructor arguments of MixingFeatureExtractor (src/mixing_utils.py:19-27)."""
    sample_rate: int = 44100
    n_fft: int = 1024
    hop_length: int = 256
    n_mels: int = 128
    use_detailed_spectral: bool = False
    n_spectral_bins: int = 32

    @property
    def n_freq(self) -> int:
        return self.n_fft // 2 + 1

    def n_frames(self, T: int) -> int:
        return 1 + T // self.hop_length                     # center=True framing of torch.stft

    @property
    def spectral_dim(self) -> int:                          # :62
        return (self.n_spectral_bins + 2) if self.use_detailed_spectral else 5

    @property
    def feature_dim(self) -> int:                           # :53-69
        return 4 * (6 + self.spectral_dim + 3) + 4 + 4


# --------------------------------------------------------------------------
# mel filterbank + window (torchaudio functional.py:518-587, _transforms.py:96)
# --------------------------------------------------------------------------
def hann_periodic(n_fft: int, dtype=torch.float32) -> torch.Tensor:
    return torch.hann_window(n_fft, periodic=True, dtype=dtype)


def htk_mel_filterbank(sample_rate: int, n_fft: int, n_mels: int) -> torch.Tensor:
    """(n_freq, n_mels) fp32 triangular HTK filterbank, f_min=0, f_max=sr//2, norm=None."""
    n_freq = n_fft // 2 + 1
    f_max = float(sample_rate // 2)
    bin_hz = torch.linspace(0, sample_rate // 2, n_freq)
    mel_lo = 2595.0 * math.log10(1.0 + 0.0 / 700.0)
    mel_hi = 2595.0 * math.log10(1.0 + f_max / 700.0)
    mel_pts = torch.linspace(mel_lo, mel_hi, n_mels + 2)
    hz_pts = 700.0 * (10.0 ** (mel_pts / 2595.0) - 1.0)
    width = hz_pts[1:] - hz_pts[:-1]
    dist = hz_pts.unsqueeze(0) - bin_hz.unsqueeze(1)        # (n_freq, n_mels+2)
    falling = (-1.0 * dist[:, :-2]) / width[:-1]
    rising = dist[:, 2:] / width[1:]
    return torch.max(torch.zeros(1), torch.min(falling, rising))


class MelFrontEnd:
    """power-mel of (..., T) audio: reflect-pad n_fft/2, periodic Hann, rfft, |.|^2, @fb."""

    def __init__(self, cfg: FrontEndConfig, dtype=torch.float32, device="cpu"):
        # device: "cpu" for the oracle proper; a CUDA device only for bench.py's torch_cuda_comparator, which times
        # this same reference math (torch.stft -> cuFFT, dense matmul -> cuBLAS) on the GPU the kernels run on
        self.cfg = cfg
        self.dtype = dtype
        self.window = hann_periodic(cfg.n_fft).to(dtype).to(device)
        self.fb = htk_mel_filterbank(cfg.sample_rate, cfg.n_fft, cfg.n_mels).to(dtype).to(device)

    def power_spectrogram(self, x: torch.Tensor) -> torch.Tensor:
        c = self.cfg
        lead = x.shape[:-1]
        flat = x.to(self.dtype).reshape(-1, x.shape[-1])
        spec = torch.stft(flat, c.n_fft, hop_length=c.hop_length, win_length=c.n_fft,
                          window=self.window, center=True, pad_mode="reflect",
                          normalized=False, onesided=True, return_complex=True)
        spec = spec.reshape(lead + spec.shape[-2:])
        return spec.abs().pow(2.0)                           # functional.py:141-144

    def mel_power(self, x: torch.Tensor) -> torch.Tensor:
        p = self.power_spectrogram(x)                        # (..., n_freq, F)
        return torch.matmul(p.transpose(-1, -2), self.fb).transpose(-1, -2)   # _transforms.py:417


def log_mel(stems: Dict[str, torch.Tensor], cfg: FrontEndConfig = FrontEndConfig(),
            dtype=torch.float32, fe: Optional[MelFrontEnd] = None) -> torch.Tensor:
    """src/model.py:41-67: dict of 4 x (B,2,T) -> (B,8,n_mels,F) natural-log mel power."""
    fe = fe or MelFrontEnd(cfg, dtype)
    per_stem = [fe.mel_power(stems[name]) for name in STEMS]
    return torch.log(torch.cat(per_stem, dim=1) + 1e-10)


# --------------------------------------------------------------------------
# mixing features (src/mixing_utils.py:71-357)
# --------------------------------------------------------------------------
def loudness(x: torch.Tensor) -> torch.Tensor:
    """:311-318 -- NOT K-weighted, NOT gated: -0.691 + 10 log10(mean(x^2) + 1e-10)."""
    rms = torch.sqrt(torch.mean(x ** 2))
    return -0.691 + 10 * torch.log10(rms ** 2 + 1e-10)


def dynamics_features(x: torch.Tensor) -> torch.Tensor:
    """:107-139 -> [rms_L, rms_R, crest_L, crest_R, loud, loud]."""
    rms = torch.sqrt(torch.mean(x ** 2, dim=-1))
    peak = x.abs().max(dim=-1).values
    crest = 20 * torch.log10(peak / (rms + 1e-8))
    ld = loudness(x)
    # the reference round-trips the loudness through a Python float (:134) -- exact
    ld2 = torch.tensor([ld.item(), ld.item()], dtype=x.dtype, device=x.device)
    return torch.cat([rms, crest, ld2])


def _pearson_vs_index(curve: torch.Tensor) -> torch.Tensor:
    """tilt: 0 when the unbiased std is < 1e-6 (:187,:222), else corrcoef with arange (:190)."""
    if curve.std() < 1e-6:
        return torch.tensor(0.0, dtype=curve.dtype, device=curve.device)
    idx = torch.arange(curve.shape[0], dtype=curve.dtype, device=curve.device)
    return torch.corrcoef(torch.stack([idx, curve]))[0, 1]


def spectral_features(x: torch.Tensor, fe: MelFrontEnd) -> torch.Tensor:
    """:141-236 (both the 3-band and the detailed mode)."""
    c = fe.cfg
    mel = fe.mel_power(x)                                        # (2, n_mels, F)
    band_db = (10 * torch.log10(mel + 1e-10)).mean(dim=(0, 2))   # (n_mels,)
    flat = torch.exp(torch.mean(torch.log(mel + 1e-10))) / (torch.mean(mel) + 1e-10)
    if not c.use_detailed_spectral:
        lo = band_db.shape[0] // 4
        hi = band_db.shape[0] // 4 * 3
        bands = [band_db[:lo].mean(), band_db[lo:hi].mean(), band_db[hi:].mean()]
        return torch.stack(bands + [_pearson_vs_index(band_db), flat]).to(x.dtype)
    if c.n_spectral_bins >= c.n_mels:
        curve = band_db
    else:
        curve = torch.nn.functional.interpolate(band_db[None, None], size=c.n_spectral_bins,
                                                mode="linear", align_corners=True)[0, 0]
    # the reference correlates against arange(n_spectral_bins) even when curve has n_mels entries
    # (:220); corrcoef would raise on a length mismatch, which only happens for n_bins > n_mels.
    if curve.std() < 1e-6:
        tilt = torch.tensor(0.0, dtype=curve.dtype, device=curve.device)
    else:
        idx = torch.arange(c.n_spectral_bins, dtype=curve.dtype, device=curve.device)
        tilt = torch.corrcoef(torch.stack([idx, curve]))[0, 1]
    return torch.cat([curve, torch.stack([tilt, flat])]).to(x.dtype)


def stereo_features(x: torch.Tensor) -> torch.Tensor:
    """:238-268 -> [ILD, L/R Pearson correlation, mid/side energy ratio]."""
    L, R = x[0], x[1]
    rl = torch.sqrt(torch.mean(L ** 2))
    rr = torch.sqrt(torch.mean(R ** 2))
    ild = 20 * torch.log10(rl / (rr + 1e-8))
    lc, rc = L - L.mean(), R - R.mean()
    corr = (lc * rc).sum() / (torch.sqrt((lc ** 2).sum() * (rc ** 2).sum()) + 1e-8)
    mid, side = (L + R) / 2, (L - R) / 2
    msr = torch.mean(side ** 2) / (torch.mean(mid ** 2) + 1e-8)
    return torch.stack([ild, corr, msr])


def masking_features(stems: Dict[str, torch.Tensor], fe: MelFrontEnd) -> torch.Tensor:
    """:270-309 -- mean_{k,t} sigmoid(max_{j!=i} S_j - S_i) on channel-mean mel POWER."""
    S = [fe.mel_power(stems[n]).mean(dim=0) for n in STEMS]
    out = []
    for i in range(4):
        others = torch.stack([S[j] for j in range(4) if j != i]).max(dim=0).values
        out.append(torch.sigmoid((0.0 - (S[i] - others)) / 1.0).mean())
    return torch.stack(out)


def feature_layout(cfg: FrontEndConfig = FrontEndConfig()) -> List[str]:
    """Names of the D entries in _flatten_features order (sorted keys, :324)."""
    sd = cfg.spectral_dim
    if cfg.use_detailed_spectral:
        spec = [f"curve{i}" for i in range(sd - 2)] + ["tilt", "flatness"]
    else:
        spec = ["low", "mid", "high", "tilt", "flatness"]
    per = (["rms_L", "rms_R", "crest_L", "crest_R", "loud", "loud_dup", "rel_loud"]
           + spec + ["ild", "corr", "msr"])
    names: List[str] = []
    for s in _SORTED_STEMS:
        if s == "other":                                       # 'masking' sorts between drums_* and other_*
            names += [f"masking_{m}" for m in STEMS]
        names += [f"{s}_{p}" for p in per]
    return names


def mixing_features(stems: Dict[str, torch.Tensor], cfg: FrontEndConfig = FrontEndConfig(),
                    dtype=torch.float32, fe: Optional[MelFrontEnd] = None) -> torch.Tensor:
    """:71-105 + :320-357 for one clip: dict of 4 x (2,T) -> (D,)."""
    fe = fe or MelFrontEnd(cfg, dtype)
    st = {n: stems[n].to(dtype) for n in STEMS}                # KeyError on a missing stem, as the reference
    mix = 0
    for n in stems:                                            # left fold in dict order (:85)
        mix = mix + stems[n].to(dtype)
    mix_loud = loudness(mix)
    blocks = {}
    for n in STEMS:
        blocks[n] = torch.cat([
            dynamics_features(st[n]),
            (loudness(st[n]) - mix_loud).reshape(1),
            spectral_features(st[n], fe),
            stereo_features(st[n]),
        ])
    masking = masking_features(st, fe)
    parts = []
    for s in _SORTED_STEMS:
        if s == "other":
            parts.append(masking)
        parts.append(blocks[s])
    vec = torch.cat(parts)
    vec = torch.clamp(vec, min=-100.0, max=100.0)              # :339
    vec = torch.where(torch.isnan(vec), torch.zeros_like(vec), vec)   # :343-348
    if torch.isinf(vec).any():                                 # :351-355 (unreachable after clamp)
        raise ValueError("Inf detected in features after clamping")
    return vec


# --------------------------------------------------------------------------
# Butterworth design + DF2T cascade (scipy.signal.butter(output='sos') / sosfilt)
# --------------------------------------------------------------------------
def butter_sos(order: int, cutoff_hz: float, btype: str, fs: float) -> np.ndarray:
    """Restates scipy.signal.butter(order, cutoff, btype, fs=fs, output='sos') for even orders.

    Analog prototype poles on the unit circle, pre-warped lp2lp / lp2hp, bilinear transform at the
    normalised rate 2, conjugate pairs turned into biquads ordered by pole radius (closest to the
    unit circle last), overall gain folded into the first section -- scipy's zpk2sos('nearest').
    """
    assert order % 2 == 0 and btype in ("low", "high")
    wn = 2.0 * float(cutoff_hz) / float(fs)
    fs2 = 2.0
    warped = 2.0 * fs2 * math.tan(math.pi * wn / fs2)
    m = np.arange(-order + 1, order, 2)
    p_analog = -np.exp(1j * np.pi * m / (2 * order))
    if btype == "low":
        p_s = warped * p_analog
        k_s = warped ** order
        z_digital = -np.ones(order)
    else:
        p_s = warped / p_analog
        k_s = float(np.real(1.0 / np.prod(-p_analog)))
        z_digital = np.ones(order)
    p_z = (2 * fs2 + p_s) / (2 * fs2 - p_s)
    if btype == "low":
        k_z = k_s * float(np.real(1.0 / np.prod(2 * fs2 - p_s)))
    else:
        k_z = k_s * float(np.real(np.prod(2 * fs2 - np.zeros(order)) / np.prod(2 * fs2 - p_s)))
    upper = sorted([p for p in p_z if p.imag > 0], key=lambda p: abs(p))   # ascending radius
    sos = np.zeros((order // 2, 6))
    zsign = -z_digital[0]                                      # (z - z0)^2 = z^2 + 2*zsign*z + 1
    for i, p in enumerate(upper):
        sos[i] = [1.0, 2.0 * zsign, 1.0, 1.0, -2.0 * p.real, abs(p) ** 2]
    sos[0, :3] *= k_z
    return sos


_LIB: Optional[ctypes.CDLL] = None
_HERE = os.path.dirname(os.path.abspath(__file__))


def build_c_helpers(force: bool = False) -> str:
    """gcc-compile oracle/oracle_dsp.c into oracle/_build/liboracle_dsp.so (checker build only)."""
    src = os.path.join(_HERE, "oracle_dsp.c")
    out_dir = os.path.join(_HERE, "_build")
    out = os.path.join(out_dir, "liboracle_dsp.so")
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        os.makedirs(out_dir, exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-o", out, src, "-lm"])
    return out


def _lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build_c_helpers())
        _LIB.oracle_sosfilt_df2t.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_long]
        _LIB.oracle_xcorr_same.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p,
                                           ctypes.c_long, ctypes.c_void_p]
    return _LIB


def sosfilt_df2t(sos: np.ndarray, x: np.ndarray) -> np.ndarray:
    """Transposed direct-form-II biquad cascade along the last axis, fp64, zero initial state."""
    sos = np.ascontiguousarray(sos, dtype=np.float64)
    x64 = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x64)
    rows = x64.reshape(-1, x64.shape[-1])
    yr = y.reshape(-1, x64.shape[-1])
    for r in range(rows.shape[0]):
        _lib().oracle_sosfilt_df2t(sos.ctypes.data, sos.shape[0], rows[r].ctypes.data,
                                   yr[r].ctypes.data, rows.shape[1])
    return y


def xcorr_same_direct(x: np.ndarray, h: np.ndarray) -> np.ndarray:
    """out[n] = sum_k x[n + k - L//2] h[k], n in [0,T): what F.conv1d(padding=L//2)[:T] computes."""
    x32 = np.ascontiguousarray(x, dtype=np.float32)
    h32 = np.ascontiguousarray(h, dtype=np.float32)
    out = np.empty_like(x32)
    _lib().oracle_xcorr_same(x32.ctypes.data, x32.shape[0], h32.ctypes.data, h32.shape[0],
                             out.ctypes.data)
    return out


def xcorr_same_fft(x: np.ndarray, h: np.ndarray) -> np.ndarray:
    """Same quantity via fp64 FFTs (for full-length clips where the direct form takes minutes)."""
    T, L = x.shape[-1], h.shape[-1]
    n = 1
    while n < T + L:
        n *= 2
    X = np.fft.rfft(np.asarray(x, dtype=np.float64), n)
    H = np.fft.rfft(np.asarray(h, dtype=np.float64), n)
    full = np.fft.irfft(X * np.conj(H), n)                     # full[m] = sum_k x[m+k] h[k] (circular)
    idx = (np.arange(T) - L // 2) % n
    return full[..., idx]


# --------------------------------------------------------------------------
# augmentation (src/mixing_utils.py:364-479) with the random decisions made explicit
# --------------------------------------------------------------------------
@dataclass
class StemAug:
    gain_db: Optional[torch.Tensor] = None       # (1,) fp32 tensor when the gain gate fired
    tilt: Optional[str] = None                   # 'high' | 'low' | None
    compress: bool = False
    cutoff_hz: Optional[float] = None            # python float (cutoff.item()) when the band-limit gate fired


@dataclass
class ClipAug:
    stems: Dict[str, StemAug] = field(default_factory=dict)
    impulse: Optional[torch.Tensor] = None       # (L,) fp32 reverb impulse when the reverb gate fired


def draw_clip_aug(prob: float = 0.5, gain_range: float = 9.0, sample_rate: int = 44100,
                  stem_names: Sequence[str] = STEMS, device="cpu") -> ClipAug:
    """Consume torch's global CPU generator in exactly the reference's (data-dependent) order."""
    clip = ClipAug()
    for name in stem_names:
        s = StemAug()
        if torch.rand(1) < prob:                                   # :389
            s.gain_db = torch.rand(1) * 2 * gain_range - gain_range
        if torch.rand(1) < prob:                                   # :395
            s.tilt = "high" if torch.rand(1) < 0.5 else "low"      # :424
        if torch.rand(1) < prob:                                   # :399
            s.compress = True
        if torch.rand(1) < prob:                                   # :403
            s.cutoff_hz = (torch.rand(1) * 8000 + 4000).item()     # :451-452
        clip.stems[name] = s
    if torch.rand(1) < prob:                                       # :409
        decay = 0.5
        L = int(sample_rate * decay)
        t = torch.linspace(0, decay, L, device=device)
        clip.impulse = torch.exp(-t / (decay / 4)) * torch.randn(L, device=device) * 0.1   # :461-463
    return clip


def compress_static(x: torch.Tensor, threshold: float = -20, ratio: float = 4) -> torch.Tensor:
    """:435-447 memoryless dB-domain compressor."""
    db = 20 * torch.log10(torch.abs(x) + 1e-8)
    over = db > threshold
    out_db = db.clone()
    out_db[over] = threshold + (db[over] - threshold) / ratio
    return torch.sign(x) * (10 ** (out_db / 20))


def reverb_standalone(x: torch.Tensor, h: torch.Tensor, fft: bool = True) -> torch.Tensor:
    """:458-479 with the impulse made explicit: 0.7 x + 0.3 conv1d(x, h, padding=L//2)[:, :T] per channel."""
    xc = xcorr_same_fft if fft else xcorr_same_direct
    hn = h.numpy()
    rev = torch.stack([torch.from_numpy(np.asarray(xc(x[c].numpy(), hn))).float() for c in range(x.shape[0])])
    return x * (1 - 0.3) + rev * 0.3


def draw_impulse(decay: float = 0.5, sample_rate: int = 44100, device="cpu") -> torch.Tensor:
    """:460-463 -- consumes torch's global generator of `device` like apply_reverb does."""
    L = int(sample_rate * decay)
    t = torch.linspace(0, decay, L, device=device)
    return torch.exp(-t / (decay / 4)) * torch.randn(L, device=device) * 0.1


def spectral_tilt(x: torch.Tensor, kind: str, sample_rate: int = 44100) -> torch.Tensor:
    """:421-433 with the high/low draw made explicit."""
    sos = butter_sos(2, 2000, "high", sample_rate) if kind == "high" else butter_sos(2, 500, "low", sample_rate)
    return _filter(x, sos)


def bandwidth_limit(x: torch.Tensor, cutoff_hz: float, sample_rate: int = 44100) -> torch.Tensor:
    """:449-456 with the cutoff draw made explicit."""
    return _filter(x, butter_sos(4, cutoff_hz, "low", sample_rate))


def song_a_clip(pcm: np.ndarray, T: int = 441000) -> Dict[str, torch.Tensor]:
    """BASELINE config 1 / SURVEY section 8d: the 4-stem clip cut from assets/song_A.wav (int16 (frames, 2)):
    vocals, bass, drums, other = the asset from 0, 1.5, 3.0 and 4.5 s."""
    x = torch.from_numpy(pcm.astype(np.float32) / 32768.0).t().contiguous()
    return {n: x[:, o:o + T].contiguous() for n, o in zip(STEMS, (0, 66150, 132300, 198450))}


def _filter(x: torch.Tensor, sos: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(sosfilt_df2t(sos, x.numpy())).float()     # fp64 inside, .float() after


def apply_clip_aug(stems: Dict[str, torch.Tensor], aug: ClipAug, sample_rate: int = 44100,
                   fft_reverb: bool = True) -> Dict[str, torch.Tensor]:
    """:372-419 with the decisions of ``aug``; dict of (2,T) fp32 -> new dict of (2,T) fp32."""
    out: Dict[str, torch.Tensor] = {}
    for name, x in stems.items():
        a = aug.stems[name]
        y = x.clone()
        if a.gain_db is not None:
            y = y * (10 ** (a.gain_db / 20))
        if a.tilt is not None:
            sos = butter_sos(2, 2000, "high", sample_rate) if a.tilt == "high" \
                else butter_sos(2, 500, "low", sample_rate)
            y = _filter(y, sos)
        if a.compress:
            y = compress_static(y)
        if a.cutoff_hz is not None:
            y = _filter(y, butter_sos(4, a.cutoff_hz, "low", sample_rate))
        out[name] = y
    if aug.impulse is not None:
        mix = sum(out.values())
        h = aug.impulse.numpy()
        xc = xcorr_same_fft if fft_reverb else xcorr_same_direct
        rev = torch.stack([torch.from_numpy(np.asarray(xc(mix[c].numpy(), h))).float()
                           for c in range(mix.shape[0])])
        wet = mix * (1 - 0.3) + rev * 0.3                              # :478-479
        total = sum(torch.mean(s ** 2) for s in out.values()) + 1e-8    # :413
        for name in out:
            share = torch.mean(out[name] ** 2) / total
            out[name] = out[name] + wet * share * 0.3                  # :417
    return out


# --------------------------------------------------------------------------
# synthetic inputs shared by tests / bench
# --------------------------------------------------------------------------
def synthetic_clip(seed: int, T: int, amp: float = 0.25) -> Dict[str, torch.Tensor]:
    """U(-amp, amp) stems drawn in order vocals,bass,drums,other from a seeded CPU generator."""
    g = torch.Generator().manual_seed(seed)
    return {n: (torch.rand(2, T, generator=g) * 2 - 1) * amp for n in STEMS}


def musical_clip(seed: int, T: int, sample_rate: int = 44100) -> Dict[str, torch.Tensor]:
    """Deterministic 'music-like' stems: decaying partials, band-limited noise, stereo spread.

    Gives a realistic spectral dynamic range (quiet high-mel bins) that white noise lacks.
    """
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(T, dtype=torch.float64) / sample_rate
    out = {}
    for si, name in enumerate(STEMS):
        f0 = [220.0, 55.0, 110.0, 330.0][si] * (1 + 0.1 * torch.rand(1, generator=g).item())
        sig = torch.zeros(2, T, dtype=torch.float64)
        for h in range(1, 9):
            ph = torch.rand(2, generator=g) * 2 * math.pi
            pan = torch.rand(2, generator=g) * 0.8 + 0.2
            amp = 0.3 / h ** (1.0 + 0.5 * si)
            env = torch.exp(-((t * (1.5 + si)) % 1.0) * (2.0 + h))
            for c in range(2):
                sig[c] += amp * pan[c] * env * torch.sin(2 * math.pi * f0 * h * t + ph[c])
        noise = torch.randn(2, T, generator=g, dtype=torch.float64)
        k = [64, 8, 2, 16][si]
        noise = torch.nn.functional.avg_pool1d(noise[None], k, 1, padding=k // 2)[0, :, :T]
        sig += 0.02 * noise * [0.5, 0.2, 2.0, 1.0][si]
        out[name] = sig.float()
    return out
