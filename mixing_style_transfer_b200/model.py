"""Drop-in for the reference's MelSpectrogramPreprocessor (src/model.py:17-67) on libmixfe.

Keeps the module path of the two persistent buffers a reference checkpoint carries
(`mel_transform.spectrogram.window`, `mel_transform.mel_scale.fb` -- SURVEY.md section 5), so
`load_state_dict(strict=True)` of the reference's encoder (src/train.py:78) keeps working when this
class replaces the reference's.
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn as nn

from ._native import STEMS, Plan
from .mixing_utils import _gather_stems, _resolve_device, hann_window_periodic, htk_mel_filterbank


class _Spectrogram(nn.Module):
    def __init__(self, n_fft: int):
        super().__init__()
        self.register_buffer("window", hann_window_periodic(n_fft))


class _MelScale(nn.Module):
    def __init__(self, sample_rate: int, n_fft: int, n_mels: int):
        super().__init__()
        self.register_buffer("fb", htk_mel_filterbank(sample_rate, n_fft, n_mels))


class _MelTransform(nn.Module):
    """Holds the buffers under the names torchaudio.transforms.MelSpectrogram uses."""

    def __init__(self, sample_rate: int, n_fft: int, n_mels: int):
        super().__init__()
        self.spectrogram = _Spectrogram(n_fft)
        self.mel_scale = _MelScale(sample_rate, n_fft, n_mels)


class MelSpectrogramPreprocessor(nn.Module):
    """Converts raw audio stems to mel-spectrograms: 4 x (B,2,T) -> (B,8,n_mels,F), natural log of
    (mel power + 1e-10), channel order [vL,vR,bL,bR,dL,dR,oL,oR] (reference: src/model.py:41-67)."""

    def __init__(self, sample_rate=44100, n_fft=1024, hop_length=256, n_mels=128):
        super().__init__()
        self.sample_rate = sample_rate
        self.n_fft = n_fft
        self.hop_length = hop_length
        self.n_mels = n_mels
        self.mel_transform = _MelTransform(sample_rate, n_fft, n_mels)
        self._plans: Dict[int, Plan] = {}
        self._fb_versions: Dict[int, tuple] = {}       # per device: the filterbank its plan was built from

    def _plan(self, device: torch.device) -> Plan:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        fb = self.mel_transform.mel_scale.fb
        version = (fb._version, fb.data_ptr())
        pl = self._plans.get(idx)
        if pl is None:
            pl = Plan(torch.device("cuda", idx), self.sample_rate, self.n_fft, self.hop_length, self.n_mels, 0, fb=fb)
            self._plans[idx] = pl
        elif version != self._fb_versions.get(idx):
            pl.set_filterbank(fb)              # a checkpoint load replaced the buffer
        self._fb_versions[idx] = version
        return pl

    def forward(self, stems_dict):
        first = stems_dict[STEMS[0]]
        src_device = first.device
        device = _resolve_device(first, None)
        plan = self._plan(device)
        if torch.is_grad_enabled() and any(stems_dict[n].requires_grad for n in STEMS):
            # the style-transfer loop differentiates through the encoder's log-mel (src/train_style_transfer.py:206-215)
            lm = _LogMelFunction.apply(plan, device, *[stems_dict[n] for n in STEMS])
        else:
            stems, _ = _gather_stems(stems_dict, device)
            _, lm = plan.features_logmel(stems, want_features=False, want_logmel=True)
        return lm if lm.device == src_device else lm.to(src_device)


class _LogMelFunction(torch.autograd.Function):
    """log-mel with a hand-written backward (kernels K6F / K6); the forward saves its inputs and its output."""

    @staticmethod
    def forward(ctx, plan, device, *stems):
        gathered, _ = _gather_stems(dict(zip(STEMS, stems)), device)
        _, lm = plan.features_logmel(gathered, want_features=False, want_logmel=True)
        ctx.plan = plan
        ctx.meta = [(x.dim(), x.device, x.dtype) for x in stems]
        ctx.save_for_backward(*gathered, lm)         # the output too: dmel = g * exp(-logmel), no mel recomputation (K6F)
        return lm

    @staticmethod
    def backward(ctx, grad_logmel):
        *saved, lm = ctx.saved_tensors
        grads = ctx.plan.logmel_backward(list(saved), grad_logmel, logmel=lm)
        out = []
        for g, (dim, dev, dt), need in zip(grads, ctx.meta, ctx.needs_input_grad[2:]):
            if not need:
                out.append(None)
                continue
            if dim == 2:
                g = g[0]
            out.append(g.to(device=dev, dtype=dt))
        return (None, None, *out)
