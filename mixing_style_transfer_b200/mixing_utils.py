"""Drop-in host-side mirror of the reference's src/mixing_utils.py, running on libmixfe (sm_100a).

Same class names, constructor signatures, attribute names, tensor layouts and error behaviour as
the reference (barry-mir/mixing-style-transfer):

* ``MixingFeatureExtractor``  -- src/mixing_utils.py:16-357
* ``AudioAugmenter``          -- src/mixing_utils.py:364-479

What differs, on purpose: everything is computed by hand-written CUDA kernels through the C ABI
of ``include/mixfe.h``; a whole batch ``(B,2,T)`` is accepted wherever the reference takes one
clip ``(2,T)``; the 4 + 4 + 4 STFTs the reference spends per clip on spectral features, masking
and the encoder's log-mel are one fused pass (``features_and_logmel``).  CPU tensors are copied
to the GPU and the result is copied back -- a transfer, not a CPU compute path; there is none.
"""
from __future__ import annotations

import ctypes
import math
from typing import Dict, List, Optional, Tuple

import torch

from . import _native
from ._native import STEMS, AugParams, Plan


def hann_window_periodic(n_fft: int) -> torch.Tensor:
    """The `window` buffer torchaudio's Spectrogram registers (torch.hann_window, periodic)."""
    return torch.hann_window(n_fft, periodic=True, dtype=torch.float32)


def htk_mel_filterbank(sample_rate: int, n_fft: int, n_mels: int) -> torch.Tensor:
    """(n_fft//2+1, n_mels) fp32 triangular filterbank with the parameters the reference's
    MelSpectrogram uses: f_min=0, f_max=sample_rate//2, HTK mel scale, no area normalisation.
    Evaluated with fp32 tensor ops so that it equals the `mel_scale.fb` buffer of a reference
    checkpoint bit for bit."""
    n_freq = n_fft // 2 + 1
    top_hz = float(sample_rate // 2)
    bins_hz = torch.linspace(0, sample_rate // 2, n_freq)
    mel_top = 2595.0 * math.log10(1.0 + top_hz / 700.0)
    mel_edges = torch.linspace(0.0, mel_top, n_mels + 2)
    hz_edges = 700.0 * (10.0 ** (mel_edges / 2595.0) - 1.0)
    spans = hz_edges[1:] - hz_edges[:-1]
    offsets = hz_edges.unsqueeze(0) - bins_hz.unsqueeze(1)
    rise = (-1.0 * offsets[:, :-2]) / spans[:-1]
    fall = offsets[:, 2:] / spans[1:]
    return torch.clamp(torch.min(rise, fall), min=0.0)


def _resolve_device(t: torch.Tensor, default: Optional[torch.device]) -> torch.device:
    if t.is_cuda:
        return t.device
    if default is not None:
        return torch.device(default)
    if not torch.cuda.is_available():
        raise RuntimeError("mixing_style_transfer_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _gather_stems(stems_dict: Dict[str, torch.Tensor], device: torch.device) -> Tuple[List[torch.Tensor], bool]:
    """dict of (2,T) or (B,2,T) -> list of 4 (B,2,T) CUDA fp32 tensors in STEMS order, was_batched."""
    out = []
    batched = None
    for name in STEMS:
        x = stems_dict[name]                         # KeyError for a missing stem, as the reference
        if x.dim() == 2:
            b = False
            x = x.unsqueeze(0)
        elif x.dim() == 3:
            b = True
        else:
            raise ValueError(f"stem '{name}' must be (2,T) or (B,2,T), got {tuple(x.shape)}")
        if batched is None:
            batched = b
        elif batched != b:
            raise ValueError("stems mix batched and unbatched tensors")
        if x.shape[1] != 2:
            raise ValueError(f"stem '{name}' must be stereo, got {tuple(x.shape)}")
        if not x.is_cuda or x.device != device:
            x = x.to(device, non_blocking=True)
        out.append(x.detach())
    return out, bool(batched)


class MixingFeatureExtractor:
    """Extract interpretable mixing features from audio stems (reference: src/mixing_utils.py:16)."""

    def __init__(self, sample_rate=44100, n_fft=1024, hop_length=256, n_mels=128,
                 use_detailed_spectral=False, n_spectral_bins=32, device=None):
        self.sr = sample_rate
        self.n_fft = n_fft
        self.hop_length = hop_length
        self.n_mels = n_mels
        self.use_detailed_spectral = use_detailed_spectral
        self.n_spectral_bins = n_spectral_bins
        self._device = torch.device(device) if device is not None else None
        self._fb = htk_mel_filterbank(sample_rate, n_fft, n_mels)
        self._plans: Dict[int, Plan] = {}
        self._plans2: Dict[tuple, Plan] = {}          # plans of `second` configurations (features_and_logmel)

    # -- reference API ---------------------------------------------------------------------
    def get_feature_dim(self):
        """src/mixing_utils.py:53-69."""
        spectral_dim = 5 if not self.use_detailed_spectral else (self.n_spectral_bins + 2)
        return 4 * (6 + spectral_dim + 3) + 4 + 4

    def extract_all_features(self, stems_dict):
        """src/mixing_utils.py:71-105: dict of 4 x (2,T) -> (D,); batched (B,2,T) -> (B,D)."""
        feats, _ = self._run(stems_dict, True, False)
        return feats

    # -- additive API ----------------------------------------------------------------------
    def features_and_logmel(self, stems_dict, second=None, chunk_clips: Optional[int] = None):
        """One fused pass: (features (D,)|(B,D), logmel (B,8,n_mels,F)) -- the encoder's
        MelSpectrogramPreprocessor output (src/model.py:41-67) for the same (n_fft, hop, n_mels).

        second = (n_fft2, hop2, n_mels2): ALSO the log-mel of a second STFT configuration -- the reference trains its
        encoder on (2048, 512, 80) while the extractor keeps its own triple (src/train_style_transfer.py:523-533,
        scripts/train_baseline.sh:42-44).  Returns (features, logmel, logmel2 (B,8,n_mels2,F2)).
        Default (`chunk_clips=None`): two whole-batch passes, the fastest (72 x 10 s clips: 3.65 ms; neither kernel is
        bound by HBM, so reading the audio twice costs nothing).  `chunk_clips=n` walks the batch n clips at a time with
        the second configuration's kernel right behind the first one's on each chunk: with 4 clips (56 MB of 10 s
        audio, inside the 126 MB L2) the second kernel's loads mostly hit L2 (1.41 GB read from
        DRAM instead of 2.07 GB, profiles/r03b_two_config_traffic.txt) at 4.76 ms -- small launches fill the GPU less
        well; the mode for a pipeline whose other kernels need the HBM bandwidth.  (Two kernels sharing the samples
        through L2, not one kernel sharing them through shared memory)"""
        if second is None:
            return self._run(stems_dict, True, True)
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, self._device)
        stems, batched = _gather_stems(stems_dict, device)
        B, _, T = stems[0].shape
        plan, plan2 = self._plan(device), self._second_plan(device, tuple(int(v) for v in second))
        feats = torch.empty((B, plan.feature_dim), dtype=torch.float32, device=device)
        lm = torch.empty((B, 8, plan.n_mels, plan.num_frames(T)), dtype=torch.float32, device=device)
        lm2 = torch.empty((B, 8, plan2.n_mels, plan2.num_frames(T)), dtype=torch.float32, device=device)
        step = B if chunk_clips is None else max(1, int(chunk_clips))
        for lo in range(0, B, step):
            hi = min(B, lo + step)
            part = [x[lo:hi] for x in stems]
            plan.features_logmel(part, True, True, features_out=feats[lo:hi], logmel_out=lm[lo:hi])
            plan2.features_logmel(part, False, True, logmel_out=lm2[lo:hi])
        if not batched:
            feats = feats[0]
        if src_device != device:
            feats, lm, lm2 = feats.to(src_device), lm.to(src_device), lm2.to(src_device)
        return feats, lm, lm2

    def _second_plan(self, device: torch.device, cfg) -> Plan:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        key = (idx,) + tuple(cfg)
        pl = self._plans2.get(key)
        if pl is None:
            n_fft2, hop2, n_mels2 = cfg
            pl = self._plans2[key] = Plan(torch.device("cuda", idx), self.sr, n_fft2, hop2, n_mels2, 0,
                                          fb=htk_mel_filterbank(self.sr, n_fft2, n_mels2))
        return pl

    def gated_loudness(self, stems_dict):
        """EXTENSION (not in the reference): ITU-R BS.1770-4 integrated loudness --
        K-weighted, gated -- of the four stems and of their mixture, (5,) or (B,5) in LUFS, order vocals, bass,
        drums, other, mixture.  The reference's own loudness (un-weighted, un-gated, src/mixing_utils.py:311-318)
        stays what the feature vector holds."""
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, None)
        stems, batched = _gather_stems(stems_dict, device)
        out = self._plan(device).gated_loudness(stems)
        if not batched:
            out = out[0]
        return out if out.device == src_device else out.to(src_device)

    def short_term_loudness_iqr(self, stems_dict):
        """EXTENSION (not in the reference's code; its proposal, p.2 section 2.1, names DR_IQR = IQR(L_short(t))): the
        inter-quartile range of the short-term loudness -- K-weighted, 3 s windows every 100 ms (EBU Tech 3341), windows
        under -70 LUFS dropped -- of the four stems and of their mixture, (5,) or (B,5) in LU, order vocals, bass,
        drums, other, mixture.  Needs 3 s <= T <= 6 min 52 s."""
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, None)
        stems, batched = _gather_stems(stems_dict, device)
        out = self._plan(device).short_term_loudness_iqr(stems)
        if not batched:
            out = out[0]
        return out if out.device == src_device else out.to(src_device)

    def band_weights(self) -> torch.Tensor:
        """(n_mels,) IEC 61672 A-weighting as a power ratio, normalised to its maximum, at the centre frequency of every
        mel band (HTK scale, f_min = 0, f_max = sr // 2): the A(k) of weighted_masking."""
        top = 2595.0 * math.log10(1.0 + float(self.sr // 2) / 700.0)
        mel = torch.linspace(0.0, top, self.n_mels + 2, dtype=torch.float64)[1:-1]
        f2 = (700.0 * (10.0 ** (mel / 2595.0) - 1.0)) ** 2
        ra = (12194.0 ** 2 * f2 ** 2) / ((f2 + 20.6 ** 2) * torch.sqrt((f2 + 107.7 ** 2) * (f2 + 737.9 ** 2)) * (f2 + 12194.0 ** 2))
        w = ra ** 2
        return (w / w.max()).float()

    def weighted_masking(self, stems_dict, beta=0.0, tau=1.0):
        """EXTENSION (not in the reference's code; its proposal, p.2 section 2.4): activity- and band-weighted inter-stem
        masking M_i = sum a_i(t) A(k) sigmoid((beta - Delta_i(k,t)) / tau) / sum a_i(t) A(k), (4,) or (B,4) in the order
        vocals, bass, drums, other.  a_i(t): soft activity gate of the frame's mel energy at the level of -60 dBFS white
        noise; A(k): band_weights().  beta = 0, tau = 1 are the constants of the reference's un-weighted feature
        (src/mixing_utils.py:300-301), which stays what the 64-vector holds."""
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, self._device)
        stems, batched = _gather_stems(stems_dict, device)
        plan = self._plan(device)
        _, lm = plan.features_logmel(stems, False, True)
        out = plan.weighted_masking(lm, stems[0].shape[-1], self.band_weights(), beta, tau)
        if not batched:
            out = out[0]
        return out if out.device == src_device else out.to(src_device)

    def feature_names(self) -> List[str]:
        """Names of the D entries in the reference's sorted-key order (src/mixing_utils.py:320-357)."""
        if self.use_detailed_spectral:
            spec = [f"spectral_curve{i}" for i in range(self.n_spectral_bins)] + ["spectral_tilt", "spectral_flatness"]
        else:
            spec = ["spectral_low", "spectral_mid", "spectral_high", "spectral_tilt", "spectral_flatness"]
        per = (["rms_L", "rms_R", "crest_L", "crest_R", "loudness", "loudness_dup", "rel_loudness"]
               + spec + ["stereo_ild", "stereo_corr", "stereo_msr"])
        names: List[str] = []
        for s in sorted(STEMS):
            if s == "other":
                names += [f"masking_{m}" for m in STEMS]
            names += [f"{s}_{p}" for p in per]
        return names

    # -- internals -------------------------------------------------------------------------
    def _new_plan(self, device: torch.device) -> Plan:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self.use_detailed_spectral and self.n_spectral_bins > self.n_mels:
            # torch.stack([arange(n_bins), curve]) raises in the reference (:220)
            raise RuntimeError("stack expects each tensor to be equal size: n_spectral_bins > n_mels")
        nb = self.n_spectral_bins if self.use_detailed_spectral else 0
        return Plan(torch.device("cuda", idx), self.sr, self.n_fft, self.hop_length, self.n_mels, nb, fb=self._fb)

    def _plan(self, device: torch.device) -> Plan:
        """The extractor's plan on `device` (one per device).  A plan is single-stream: its workspace is shared by
        every call, so concurrent use from several CUDA streams needs one extractor (or plan) per stream."""
        idx = device.index if device.index is not None else torch.cuda.current_device()
        pl = self._plans.get(idx)
        if pl is None:
            pl = self._plans[idx] = self._new_plan(device)
        return pl

    def _run(self, stems_dict, want_features: bool, want_logmel: bool):
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, self._device)
        stems, batched = _gather_stems(stems_dict, device)
        feats, lm = self._plan(device).features_logmel(stems, want_features, want_logmel)
        if feats is not None:
            if not batched:
                feats = feats[0]
            if feats.device != src_device:
                feats = feats.to(src_device)
        if lm is not None and lm.device != src_device:
            lm = lm.to(src_device)
        return feats, lm


# =============================================================================================
# Augmentation
# =============================================================================================
def butter_sos(order: int, cutoff_hz: float, btype: str, fs: float) -> List[List[float]]:
    """Second-order sections of a digital Butterworth filter -- what the reference obtains from
    scipy.signal.butter(order, cutoff, btype=..., fs=fs, output='sos') at src/mixing_utils.py:426-429,452.
    Even orders, 'low' / 'high' only.  Design runs on the host in double precision (a handful of
    flops per clip); the filtering itself is the CUDA scan kernel.

    Method: Butterworth poles of the analog prototype, frequency pre-warping, bilinear transform,
    conjugate pole pairs -> biquads ordered by increasing pole radius, gain in the first section.
    """
    if order % 2 or btype not in ("low", "high"):
        raise ValueError("butter_sos: even order and btype in {'low','high'} only")
    wn = 2.0 * float(cutoff_hz) / float(fs)
    if not 0.0 < wn < 1.0:
        raise ValueError("Digital filter critical frequencies must be 0 < Wn < 1")
    k_warp = 4.0 * math.tan(math.pi * wn / 2.0)            # bilinear transform at fs = 2
    import cmath
    poles_z = []
    gain_den = 1.0 + 0.0j
    proto_prod = 1.0 + 0.0j
    for m in range(-order + 1, order, 2):
        p = -cmath.exp(1j * math.pi * m / (2 * order))     # unit-circle prototype pole
        proto_prod *= -p
        ps = k_warp * p if btype == "low" else k_warp / p
        poles_z.append((4.0 + ps) / (4.0 - ps))
        gain_den *= (4.0 - ps)
    if btype == "low":
        gain = (k_warp ** order) * (1.0 / gain_den).real
        zsign = 1.0                                        # zeros at z = -1: (z + 1)^2
    else:
        gain = (1.0 / proto_prod).real * ((4.0 ** order) / gain_den).real
        zsign = -1.0                                       # zeros at z = +1: (z - 1)^2
    upper = sorted([p for p in poles_z if p.imag > 0], key=abs)
    sos = [[1.0, 2.0 * zsign, 1.0, 1.0, -2.0 * p.real, abs(p) ** 2] for p in upper]
    for i in range(3):
        sos[0][i] *= gain
    return sos


def butter_sos_batch(order: int, cutoff_hz, btype: str, fs: float):
    """butter_sos for an array of cutoffs at once (numpy): (n, order // 2, 6).  Same arithmetic, vectorised; used by
    AudioAugmenter.draw_batch so that a batch's filter design costs one numpy pass."""
    import numpy as np
    if order % 2 or btype not in ("low", "high"):
        raise ValueError("butter_sos: even order and btype in {'low','high'} only")
    wn = 2.0 * np.asarray(cutoff_hz, dtype=np.float64).reshape(-1) / float(fs)
    if wn.size and not ((wn > 0.0) & (wn < 1.0)).all():
        raise ValueError("Digital filter critical frequencies must be 0 < Wn < 1")
    k_warp = 4.0 * np.tan(np.pi * wn / 2.0)                                   # (n,)
    m = np.arange(-order + 1, order, 2)
    p = -np.exp(1j * np.pi * m / (2 * order))                                 # (order,) prototype poles
    ps = k_warp[:, None] * p[None, :] if btype == "low" else k_warp[:, None] / p[None, :]
    pz = (4.0 + ps) / (4.0 - ps)
    gain_den = np.prod(4.0 - ps, axis=1)
    if btype == "low":
        gain = (k_warp ** order) * (1.0 / gain_den).real
        zsign = 1.0
    else:
        gain = (1.0 / np.prod(-p)).real * ((4.0 ** order) / gain_den).real
        zsign = -1.0
    sos = np.zeros((wn.size, order // 2, 6))
    if wn.size == 0:
        return sos
    # The bilinear map keeps a pole's half plane and orders the radii by the prototype's real parts, so the columns of the
    # upper-half-plane poles and their order by radius are the same for every cutoff: take them from the first row ...
    cols = np.nonzero(pz[0].imag > 0)[0]
    cols = cols[np.argsort(np.abs(pz[0, cols]))]
    q = pz[:, cols]
    rad = np.abs(q)
    if cols.size != order // 2 or not ((pz[:, cols].imag > 0).all() and (np.diff(rad, axis=1) > 0).all()):
        for i in range(wn.size):                         # ... and fall back to the per-row sort if that ever fails
            upper = sorted([v for v in pz[i] if v.imag > 0], key=abs)
            for j, v in enumerate(upper):
                sos[i, j] = [1.0, 2.0 * zsign, 1.0, 1.0, -2.0 * v.real, abs(v) ** 2]
            sos[i, 0, :3] *= gain[i]
        return sos
    sos[:, :, 0] = 1.0
    sos[:, :, 1] = 2.0 * zsign
    sos[:, :, 2] = 1.0
    sos[:, :, 3] = 1.0
    sos[:, :, 4] = -2.0 * q.real
    sos[:, :, 5] = rad ** 2
    sos[:, 0, :3] *= gain[:, None]
    return sos


class ClipAugmentation:
    """The random decisions augment_stems makes for one clip, drawn in the reference's order."""
    __slots__ = ("gain_db", "tilt", "compress", "cutoff_hz", "impulse", "comp_threshold", "comp_ratio")

    def __init__(self):
        self.gain_db: Dict[str, Optional[torch.Tensor]] = {}
        self.tilt: Dict[str, Optional[str]] = {}
        self.compress: Dict[str, bool] = {}
        self.cutoff_hz: Dict[str, Optional[float]] = {}
        self.impulse: Optional[torch.Tensor] = None
        self.comp_threshold: float = -20.0           # apply_compression's defaults (:435), what augment_stems uses
        self.comp_ratio: float = 4.0


class BatchAugmentation:
    """Decisions for n clips drawn in ONE pass (AudioAugmenter.draw_batch): packed mixfe_aug_params + impulses."""
    __slots__ = ("n", "params", "impulse", "reverb_rows", "arrays")

    def __init__(self, n, params, impulse, reverb_rows, arrays):
        self.n, self.params, self.impulse, self.reverb_rows, self.arrays = n, params, impulse, reverb_rows, arrays

    def __len__(self):
        return self.n

    def slice(self, lo: int, hi: int) -> "BatchAugmentation":
        """Decisions of clips [lo, hi) (views: no copy of the packed parameters or the impulses)."""
        import numpy as np
        n = hi - lo
        params = (AugParams * n).from_buffer(self.params, lo * ctypes.sizeof(AugParams))
        rows = self.reverb_rows[(self.reverb_rows >= lo) & (self.reverb_rows < hi)] - lo
        imp = self.impulse[lo:hi] if (self.impulse is not None and rows.size) else None
        return BatchAugmentation(n, params, imp, rows, {k: v[lo:hi] for k, v in self.arrays.items()})


def _aug_dtype():
    import numpy as np
    return np.dtype([("gain_lin", np.float32, (4,)), ("tilt_on", np.int32, (4,)), ("tilt_sos", np.float64, (4, 6)),
                     ("comp_on", np.int32, (4,)), ("comp_threshold_db", np.float32, (4,)), ("comp_ratio", np.float32, (4,)),
                     ("bw_on", np.int32, (4,)), ("bw_sos", np.float64, (4, 12)), ("reverb_on", np.int32),
                     ("_pad", np.int32)], align=True)


class AudioAugmenter:
    """Apply mixing augmentations to create different mixing styles (reference: src/mixing_utils.py:364)."""

    def __init__(self, sample_rate=44100, gain_range=9.0, prob=0.5, device=None):
        self.sr = sample_rate
        self.gain_range = gain_range
        self.prob = prob
        self._device = torch.device(device) if device is not None else None
        self._plans: Dict[int, Plan] = {}
        self._tilt_sos = {k: butter_sos(2, fc, k, self.sr)[0] for k, fc in (("high", 2000), ("low", 500))}   # :426-429
        self._env: Dict[tuple, torch.Tensor] = {}

    # -- decisions (host RNG, identical draw order to the reference) ---------------------------
    def draw(self, stem_names=STEMS, impulse_device="cpu") -> ClipAugmentation:
        """Consume torch's global CPU generator exactly as augment_stems does (:385-409, :424, :451, :461-463).
        The impulse noise is drawn on `impulse_device` -- the reference draws it on the audio's device."""
        d = ClipAugmentation()
        for name in stem_names:
            d.gain_db[name] = None
            d.tilt[name] = None
            d.compress[name] = False
            d.cutoff_hz[name] = None
            if torch.rand(1) < self.prob:
                d.gain_db[name] = torch.rand(1) * 2 * self.gain_range - self.gain_range
            if torch.rand(1) < self.prob:
                d.tilt[name] = "high" if torch.rand(1) < 0.5 else "low"
            if torch.rand(1) < self.prob:
                d.compress[name] = True
            if torch.rand(1) < self.prob:
                d.cutoff_hz[name] = (torch.rand(1) * 8000 + 4000).item()
        if torch.rand(1) < self.prob:
            d.impulse = self._draw_impulse(0.5, impulse_device)
        return d

    def _draw_impulse(self, decay: float, device) -> torch.Tensor:
        """:460-463 -- exp(-t / (decay/4)) * randn * 0.1 with t = linspace(0, decay, int(sr*decay)), on `device`."""
        n = int(self.sr * decay)
        t = torch.linspace(0, decay, n, device=device)
        return torch.exp(-t / (decay / 4)) * torch.randn(n, device=device) * 0.1

    def draw_batch(self, n: int, impulse_device="cpu") -> BatchAugmentation:
        """Decisions for n clips from ONE torch.rand call (global CPU generator) and one randn for the impulses.

        ADDITIVE API.  Same distributions and gates as augment_stems (:385-409), but the uniforms are consumed as a
        block u[n, 4, 7] = (gain gate, gain, tilt gate, tilt kind, compressor gate, band-limit gate, cutoff) followed
        by v[n] (reverb gate) instead of the reference's data-dependent sequence, so the stream differs from n calls
        of draw(); use draw() when the reference's RNG stream matters.  The filter design runs vectorised in numpy;
        the result is already packed for the kernels (no per-clip Python)."""
        import numpy as np
        u = torch.rand(n, 4, 7).numpy()
        v = torch.rand(n).numpy()
        arr = (AugParams * n)()
        a = np.frombuffer(arr, dtype=_aug_dtype())
        gain_on = u[:, :, 0] < self.prob
        gain_db = (torch.from_numpy(u[:, :, 1].copy()) * 2 * self.gain_range - self.gain_range)
        a["gain_lin"] = np.where(gain_on, (10 ** (gain_db / 20)).numpy(), np.float32(1.0))
        tilt_on = u[:, :, 2] < self.prob
        high = u[:, :, 3] < 0.5
        a["tilt_on"] = tilt_on
        hi, lo = np.asarray(self._tilt_sos["high"]), np.asarray(self._tilt_sos["low"])
        a["tilt_sos"] = np.where(tilt_on[..., None], np.where(high[..., None], hi, lo), 0.0)
        a["comp_on"] = u[:, :, 4] < self.prob
        a["comp_threshold_db"] = -20.0
        a["comp_ratio"] = 4.0
        bw_on = u[:, :, 5] < self.prob
        a["bw_on"] = bw_on
        fc = (torch.from_numpy(u[:, :, 6].copy()) * 8000 + 4000).numpy().astype(np.float64)
        if bw_on.any():
            sos = butter_sos_batch(4, fc[bw_on], "low", self.sr).reshape(-1, 12)
            full = np.zeros((n, 4, 12))
            full[bw_on] = sos
            a["bw_sos"] = full
        rev = v < self.prob
        a["reverb_on"] = rev
        rows = np.nonzero(rev)[0]
        impulse = None
        if rows.size:
            L = int(self.sr * 0.5)
            key = (L, str(impulse_device))
            env = self._env.get(key)
            if env is None:
                t = torch.linspace(0, 0.5, L, device=impulse_device)
                env = self._env[key] = torch.exp(-t / (0.5 / 4))
            # noise for every row (the kernels read row b only when reverb_on[b]): no host-dependent indexing, so the
            # draw stays asynchronous when impulse_device is the GPU
            impulse = env * torch.randn(n, L, device=impulse_device) * 0.1
        return BatchAugmentation(n, arr, impulse, rows, {"gain_on": gain_on, "gain_db": gain_db.numpy(), "tilt_on": tilt_on,
                                                         "tilt_high": high, "comp_on": np.asarray(a["comp_on"]).astype(bool),
                                                         "bw_on": bw_on, "cutoff_hz": fc, "reverb_on": rev})

    def _pack(self, decisions: List[ClipAugmentation]):
        arr = (AugParams * len(decisions))()
        for b, d in enumerate(decisions):
            p = arr[b]
            for i, name in enumerate(STEMS):
                g = d.gain_db.get(name)
                # `audio * (10 ** (gain_db / 20))` with gain_db a (1,) fp32 tensor (:390-392)
                p.gain_lin[i] = float((10 ** (g / 20)).item()) if g is not None else 1.0
                tilt = d.tilt.get(name)
                p.tilt_on[i] = 1 if tilt else 0
                if tilt:
                    sos = self._tilt_sos[tilt]
                    for j in range(6):
                        p.tilt_sos[i][j] = sos[j]
                p.comp_on[i] = 1 if d.compress.get(name) else 0
                p.comp_threshold_db[i] = float(d.comp_threshold)
                p.comp_ratio[i] = float(d.comp_ratio)
                fc = d.cutoff_hz.get(name)
                p.bw_on[i] = 1 if fc is not None else 0
                if fc is not None:
                    sos = butter_sos(4, fc, "low", self.sr)
                    for j in range(12):
                        p.bw_sos[i][j] = sos[j // 6][j % 6]
            p.reverb_on = 1 if d.impulse is not None else 0
        return arr

    def _plan(self, device: torch.device) -> Plan:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        pl = self._plans.get(idx)
        if pl is None:
            pl = Plan(torch.device("cuda", idx), self.sr)          # augmentation-only plan: no mel / STFT tables
            self._plans[idx] = pl
        return pl

    # -- reference API ---------------------------------------------------------------------
    def augment_stems(self, stems_dict):
        """src/mixing_utils.py:372-419.  dict of (2,T) -> new dict of (2,T); for (B,2,T) inputs every clip
        draws its own decisions, clip 0 first."""
        first = stems_dict[STEMS[0]]
        B = first.shape[0] if first.dim() == 3 else 1
        device = _resolve_device(first, self._device)
        imp_dev = first.device if first.is_cuda else "cpu"
        decisions = [self.draw(tuple(stems_dict.keys()), imp_dev) for _ in range(B)]
        return self.apply(stems_dict, decisions)

    def apply(self, stems_dict, decisions, out=None):
        """Run the augmentation kernels with explicit decisions: a list of ClipAugmentation (one per clip) or the
        BatchAugmentation of draw_batch.  `out`: optional dict of preallocated CUDA tensors (same shapes and
        strides as the inputs, e.g. the negatives' slice of the batch the anchors live in) that receive the
        augmented stems -- no extra copy."""
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, self._device)
        stems, batched = _gather_stems(stems_dict, device)
        B, _, T = stems[0].shape
        if len(decisions) != B:
            raise ValueError(f"{len(decisions)} decisions for {B} clips")
        impulse = None
        if isinstance(decisions, BatchAugmentation):
            params = decisions.params
            if decisions.impulse is not None:
                impulse = decisions.impulse.to(device, non_blocking=True)
        else:
            params = self._pack(decisions)
            lens = {int(d.impulse.shape[0]) for d in decisions if d.impulse is not None}
            if lens:
                if len(lens) != 1:
                    raise ValueError("impulses of different lengths in one batch")
                L = lens.pop()
                impulse = torch.zeros((B, L), dtype=torch.float32, device=device)
                for b, d in enumerate(decisions):
                    if d.impulse is not None:
                        impulse[b].copy_(d.impulse.to(device, torch.float32), non_blocking=True)
        pre = None
        if out is not None:
            pre = [out[n] if out[n].dim() == 3 else out[n].unsqueeze(0) for n in STEMS]
        outs = self._plan(device).augment(stems, params, impulse, pre)
        if out is not None:
            return {k: out[k] for k in stems_dict.keys()}
        result = {}
        for name, y in zip(STEMS, outs):
            if not batched:
                y = y[0]
            result[name] = y if y.device == src_device else y.to(src_device)
        # keep the caller's key order (the reference builds the new dict in stems_dict order)
        return {k: result[k] for k in stems_dict.keys()}

    # -- the four single-effect methods of the reference: one (C,T) tensor, or (B,C,T) ------------------------
    def _one(self, audio: torch.Tensor):
        """(C,T) | (B,C,T) on any device -> ((B,C,T) CUDA fp32, was_batched, source device)."""
        if audio.dim() not in (2, 3):
            raise ValueError(f"audio must be (C,T) or (B,C,T), got {tuple(audio.shape)}")
        device = _resolve_device(audio, self._device)
        x = audio.detach()
        if x.device != device:
            x = x.to(device, non_blocking=True)
        return (x if x.dim() == 3 else x.unsqueeze(0)), audio.dim() == 3, audio.device, device

    def _finish(self, y: torch.Tensor, batched: bool, src_device) -> torch.Tensor:
        if not batched:
            y = y[0]
        return y if y.device == src_device else y.to(src_device)

    def _effect(self, audio, fill):
        x, batched, src, device = self._one(audio)
        arr = (AugParams * x.shape[0])()
        for b in range(x.shape[0]):
            arr[b].gain_lin[0] = 1.0
            arr[b].comp_threshold_db[0] = -20.0
            arr[b].comp_ratio[0] = 4.0
            fill(arr[b])
        return self._finish(self._plan(device).effects(x, arr), batched, src)

    def apply_spectral_tilt(self, audio):
        """:421-433 -- draws the high/low decision from the global RNG like the reference (once per call)."""
        sos = self._tilt_sos["high" if torch.rand(1) < 0.5 else "low"]

        def fill(p):
            p.tilt_on[0] = 1
            for j in range(6):
                p.tilt_sos[0][j] = sos[j]
        return self._effect(audio, fill)

    def apply_compression(self, audio, threshold=-20, ratio=4):
        """:435-447 -- memoryless dB-domain compressor with the reference's two arguments."""
        if not ratio > 0:
            raise ValueError(f"ratio must be positive, got {ratio}")

        def fill(p):
            p.comp_on[0] = 1
            p.comp_threshold_db[0] = float(threshold)
            p.comp_ratio[0] = float(ratio)
        return self._effect(audio, fill)

    def apply_bandwidth_limit(self, audio):
        """:449-456 -- one cutoff drawn from the global RNG per call, like the reference."""
        sos = butter_sos(4, (torch.rand(1) * 8000 + 4000).item(), "low", self.sr)

        def fill(p):
            p.bw_on[0] = 1
            for j in range(12):
                p.bw_sos[0][j] = sos[j // 6][j % 6]
        return self._effect(audio, fill)

    def apply_envelope_compression(self, audio, threshold=-20, ratio=4, attack_ms=5.0, release_ms=50.0):
        """EXTENSION (not in the reference, whose compressor is memoryless, :435-447): the same static curve driven by an
        envelope follower -- instant-attack / exponential-release peak detector followed by a one-pole attack smoother
        (decoupled smooth peak detector, Giannoulis et al., JAES 2012), time constants in milliseconds; both 0 gives
        apply_compression.  One (C,T) tensor or a (B,C,T) batch; every row has its own detector."""
        if not ratio > 0:
            raise ValueError(f"ratio must be positive, got {ratio}")
        coeff = [math.exp(-1.0 / (ms * 1e-3 * self.sr)) if ms > 0 else 0.0 for ms in (attack_ms, release_ms)]
        x, batched, src, device = self._one(audio)
        return self._finish(self._plan(device).envelope_compress(x, threshold, ratio, coeff[0], coeff[1]), batched, src)

    def apply_reverb(self, audio, decay=0.5, impulse: Optional[torch.Tensor] = None):
        """:458-479 -- 0.7 * audio + 0.3 * conv1d(audio, impulse, padding=L//2)[:, :T] per channel, with
        L = int(sr * decay) taps drawn on the audio's device like the reference (one impulse per call, shared by the
        channels; for a (B,C,T) batch one impulse per clip, clip 0 first).  `impulse` (additive): use these taps
        ((L,) or (B,L)) instead of drawing."""
        x, batched, src, device = self._one(audio)
        B, C, T = x.shape
        if impulse is None:
            imp_dev = audio.device
            rows = [self._draw_impulse(decay, imp_dev) for _ in range(B)]
            impulse = rows[0] if B == 1 else torch.stack(rows)
        plan = self._plan(device)
        if C <= 2:
            y = plan.reverb(x, impulse)
        else:                                        # the kernel packs two channels per transform
            y = torch.cat([plan.reverb(x[:, c:c + 2], impulse) for c in range(0, C, 2)], dim=1)
        return self._finish(y, batched, src)
