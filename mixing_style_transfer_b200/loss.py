"""Drop-in for the reference's MultiResolutionSTFTLoss (src/loss.py:332-448) on libmixfe (kernel K8).

Same constructor and call: `MultiResolutionSTFTLoss(fft_sizes, hop_sizes, win_sizes, window)(x, y)` with x the
prediction and y the target, (B,C,T) or (C,T); returns the scalar mean over resolutions of spectral convergence +
log-magnitude L1, differentiable with respect to x (the cycle loss of src/train_style_transfer.py:239).  The target
gets no gradient (the reference's callers pass the input batch there).  Restrictions of the kernels: win == n_fft,
Hann window, power-of-two n_fft in [64, 4096]; anything else raises NotImplementedError.  No CPU fallback.
x and y share one packed transform, so x == y gives a loss at fp32 rounding level (~1e-7) rather than exactly 0.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn as nn

from ._native import Plan
from .mixing_utils import _resolve_device


class _STFTLossFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, module, x, y):
        device = _resolve_device(x, None)
        xd = x.detach().to(device=device, dtype=torch.float32)
        yd = y.detach().to(device=device, dtype=torch.float32)
        if xd.dim() == 2:
            xd, yd = xd.unsqueeze(0), yd.unsqueeze(0)
        if xd.shape != yd.shape or xd.dim() != 3:
            raise ValueError(f"x and y must both be (B,C,T) or (C,T), got {tuple(x.shape)} and {tuple(y.shape)}")
        plans = module._plans_for(device)
        states = [pl.stft_loss_forward(xd, yd) for pl in plans]
        total = torch.stack([s[3] + s[4] for s in states]).sum() / len(plans)
        ctx.module, ctx.plans, ctx.states = module, plans, states
        ctx.meta = (x.dim(), x.device, x.dtype)
        ctx.save_for_backward(xd, yd)
        return total.to(torch.float32).to(x.device)

    @staticmethod
    def backward(ctx, grad_out):
        xd, yd = ctx.saved_tensors
        gx = None
        if ctx.needs_input_grad[1]:
            scale = (grad_out.to(device=xd.device, dtype=torch.float32) / len(ctx.plans)).reshape(1).contiguous()
            gx = torch.zeros_like(xd, memory_format=torch.contiguous_format)
            for pl, st in zip(ctx.plans, ctx.states):
                pl.stft_loss_backward(xd, yd, st, scale, gx)
            dim, dev, dt = ctx.meta
            if dim == 2:
                gx = gx[0]
            gx = gx.to(device=dev, dtype=dt)
        if ctx.needs_input_grad[2]:
            raise NotImplementedError("MultiResolutionSTFTLoss (libmixfe): no gradient with respect to the target y")
        return None, gx, None


class MultiResolutionSTFTLoss(nn.Module):
    def __init__(self, fft_sizes: Sequence[int] = (1024, 2048, 512), hop_sizes: Sequence[int] = (256, 512, 128),
                 win_sizes: Sequence[int] = (1024, 2048, 512), window: str = "hann", sample_rate: int = 44100):
        super().__init__()
        self.fft_sizes, self.hop_sizes, self.win_sizes = list(fft_sizes), list(hop_sizes), list(win_sizes)
        self.window = window
        self.sample_rate = sample_rate
        if window != "hann":
            raise NotImplementedError("libmixfe's STFT loss uses the Hann window (the only one the reference's stft builds, :369)")
        if not (len(self.fft_sizes) == len(self.hop_sizes) == len(self.win_sizes)) or not self.fft_sizes:
            raise ValueError("fft_sizes, hop_sizes and win_sizes must have the same non-zero length")
        for n, w in zip(self.fft_sizes, self.win_sizes):
            if n != w:
                raise NotImplementedError(f"win_size {w} != fft_size {n}: libmixfe's STFT kernels need win == n_fft")
        self._plans: Dict[int, List[Plan]] = {}

    def _plans_for(self, device: torch.device) -> Tuple[Plan, ...]:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if idx not in self._plans:
            self._plans[idx] = [Plan(torch.device("cuda", idx), self.sample_rate, n, h, 64, 0)
                                for n, h in zip(self.fft_sizes, self.hop_sizes)]
        return self._plans[idx]

    def forward(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        return _STFTLossFunction.apply(self, x, y)
