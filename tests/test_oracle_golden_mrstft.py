"""CPU: the oracle of the multi-resolution STFT loss (oracle/mrstft_oracle.py) against outputs of the unmodified
reference class (tests/golden/mrstft_outputs.npz, made by tests/golden/make_golden_mrstft.py)."""
import os

import numpy as np
import torch

from oracle import mrstft_oracle as M

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mrstft_outputs.npz"))
CASES = [(1, 2, 8, 20000), (2, 1, 2, 9001), (3, 3, 8, 4500)]


def test_oracle_loss_and_gradient_equal_the_reference():
    for seed, B, C, T in CASES:
        x, y = M.pair(seed, B, C, T)
        x.requires_grad_(True)
        loss = M.mrstft_loss(x, y)
        loss.backward()
        assert abs(loss.item() - float(G[f"loss/{seed}/{B}/{C}/{T}"][0])) <= 1e-6 * abs(loss.item())
        ref = torch.from_numpy(G[f"grad/{seed}/{B}/{C}/{T}/sub7"])
        assert torch.allclose(x.grad[..., ::7], ref, rtol=1e-5, atol=1e-7 * float(G[f"grad/{seed}/{B}/{C}/{T}/absmax"][0]) + 1e-12)


def test_single_resolution():
    x, y = M.pair(9, 1, 2, 3000)
    got = M.mrstft_loss(x, y, (256,), (64,), (256,)).item()
    assert abs(got - float(G["loss_single256/9/1/2/3000"][0])) <= 1e-6 * abs(got)
