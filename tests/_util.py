import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_outputs.npz")


def golden():
    return np.load(GOLDEN)


def feature_tol(ref: torch.Tensor) -> torch.Tensor:
    """|d| <= 1e-4 |ref| + 1e-5 for every feature."""
    return 1e-4 * ref.abs() + 1e-5


def assert_features_close(got: torch.Tensor, ref: torch.Tensor, names=None, what=""):
    got, ref = got.detach().cpu().double(), ref.detach().cpu().double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    d = (got - ref).abs()
    bad = torch.nonzero(d > feature_tol(ref)).flatten().tolist()
    if bad:
        lines = [f"{i}:{names[i] if names else ''} got {got.flatten()[i]:.7g} ref {ref.flatten()[i]:.7g}" for i in bad[:12]]
        raise AssertionError(f"{what}: {len(bad)} features out of tolerance (max abs err {float(d.max()):.3g})\n" + "\n".join(lines))
    return float(d.max())


def assert_melpower_close(got: torch.Tensor, ref: torch.Tensor, what=""):
    """|d| <= 1e-4 |ref| + 1e-6 * (max over mel bins of the same frame)."""
    got, ref = got.detach().cpu().double(), ref.detach().cpu().double()
    tol = 1e-4 * ref.abs() + 1e-6 * ref.amax(dim=-2, keepdim=True) + 1e-30
    d = (got - ref).abs()
    n_bad = int((d > tol).sum())
    assert n_bad == 0, f"{what}: {n_bad} mel-power entries out of tolerance, worst ratio {float((d / tol).max()):.3g}"
    return float((d / tol).max())


def assert_logmel_close(got: torch.Tensor, ref64: torch.Tensor, what="", loose=5e-3, tight=2e-4, window=9.0):
    """Log-mel against the fp64 oracle: |d| <= `loose` everywhere, and |d| <= `tight` on every bin within `window`
    nepers of its frame's maximum (quiet bins next to loud ones are where the fp32 reference itself is only good to
    ~1e-3).  Returns (max abs, max abs on the loud bins)."""
    got, ref = got.detach().cpu().double(), ref64.detach().cpu().double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    d = (got - ref).abs()
    loud = ref >= ref.amax(dim=-2, keepdim=True) - window
    worst, worst_loud = float(d.max()), float(d[loud].max())
    assert worst <= loose, f"{what}: log-mel max abs error {worst:.3g} > {loose}"
    assert worst_loud <= tight, f"{what}: log-mel error {worst_loud:.3g} > {tight} on bins within {window} Np of the frame maximum"
    return worst, worst_loud
