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
    def features_and_logmel(self, stems_dict):
        """One fused pass: (features (D,)|(B,D), logmel (B,8,n_mels,F)) -- the encoder's
        MelSpectrogramPreprocessor output (src/model.py:41-67) for the same (n_fft, hop, n_mels)."""
        return self._run(stems_dict, True, True)

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


class ClipAugmentation:
    """The random decisions augment_stems makes for one clip, drawn in the reference's order."""
    __slots__ = ("gain_db", "tilt", "compress", "cutoff_hz", "impulse")

    def __init__(self):
        self.gain_db: Dict[str, Optional[torch.Tensor]] = {}
        self.tilt: Dict[str, Optional[str]] = {}
        self.compress: Dict[str, bool] = {}
        self.cutoff_hz: Dict[str, Optional[float]] = {}
        self.impulse: Optional[torch.Tensor] = None


class AudioAugmenter:
    """Apply mixing augmentations to create different mixing styles (reference: src/mixing_utils.py:364)."""

    def __init__(self, sample_rate=44100, gain_range=9.0, prob=0.5, device=None):
        self.sr = sample_rate
        self.gain_range = gain_range
        self.prob = prob
        self._device = torch.device(device) if device is not None else None
        self._plans: Dict[int, Plan] = {}

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
            decay = 0.5
            n = int(self.sr * decay)
            t = torch.linspace(0, decay, n, device=impulse_device)
            d.impulse = torch.exp(-t / (decay / 4)) * torch.randn(n, device=impulse_device) * 0.1
        return d

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
                    sos = butter_sos(2, 2000, "high", self.sr) if tilt == "high" else butter_sos(2, 500, "low", self.sr)
                    for j in range(6):
                        p.tilt_sos[i][j] = sos[0][j]
                p.comp_on[i] = 1 if d.compress.get(name) else 0
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
            pl = Plan(torch.device("cuda", idx), self.sr, 1024, 256, 128, 0)
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

    def apply(self, stems_dict, decisions: List[ClipAugmentation], out=None):
        """Run the augmentation kernels with explicit decisions (one per clip).  `out`: optional dict of
        preallocated CUDA tensors (same shapes and strides as the inputs, e.g. the negatives' slice of the
        batch the anchors live in) that receive the augmented stems -- no extra copy."""
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, self._device)
        stems, batched = _gather_stems(stems_dict, device)
        B, _, T = stems[0].shape
        if len(decisions) != B:
            raise ValueError(f"{len(decisions)} decisions for {B} clips")
        impulse = None
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
        outs = self._plan(device).augment(stems, self._pack(decisions), impulse, pre)
        if out is not None:
            return {k: out[k] for k in stems_dict.keys()}
        result = {}
        for name, y in zip(STEMS, outs):
            if not batched:
                y = y[0]
            result[name] = y if y.device == src_device else y.to(src_device)
        # keep the caller's key order (the reference builds the new dict in stems_dict order)
        return {k: result[k] for k in stems_dict.keys()}

    # -- the four single-effect methods of the reference (each on one (2,T) / (B,2,T) tensor) ------
    def _single(self, audio: torch.Tensor, fill):
        d = ClipAugmentation()
        for n in STEMS:
            d.gain_db[n] = None; d.tilt[n] = None; d.compress[n] = False; d.cutoff_hz[n] = None
        fill(d)
        B = audio.shape[0] if audio.dim() == 3 else 1
        zeros = torch.zeros_like(audio)
        stems = {"vocals": audio, "bass": zeros, "drums": zeros, "other": zeros}
        return self.apply(stems, [d] * B)["vocals"]

    def apply_spectral_tilt(self, audio):
        """:421-433 -- draws the high/low decision from the global RNG like the reference."""
        kind = "high" if torch.rand(1) < 0.5 else "low"
        return self._single(audio, lambda d: d.tilt.__setitem__("vocals", kind))

    def apply_compression(self, audio, threshold=-20, ratio=4):
        """:435-447 (the kernel implements the reference's fixed threshold=-20 dB, ratio=4)."""
        if threshold != -20 or ratio != 4:
            raise NotImplementedError("libmixfe's compressor is fixed at threshold=-20 dB, ratio=4 (the values "
                                      "augment_stems uses)")
        return self._single(audio, lambda d: d.compress.__setitem__("vocals", True))

    def apply_bandwidth_limit(self, audio):
        """:449-456."""
        fc = (torch.rand(1) * 8000 + 4000).item()
        return self._single(audio, lambda d: d.cutoff_hz.__setitem__("vocals", fc))

    def apply_reverb(self, audio, decay=0.5):
        """:458-479."""
        raise NotImplementedError("standalone apply_reverb: use augment_stems / apply (the reverb kernel works on "
                                  "the mixture and redistributes by stem energy, :409-417)")
