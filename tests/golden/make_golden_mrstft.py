"""Generate tests/golden/mrstft_outputs.npz from the UNMODIFIED reference loss (build container only):
values and (sub-sampled) gradients of /root/reference/src/loss.py::MultiResolutionSTFTLoss on seeded pairs
(oracle.mrstft_oracle.pair).  Run:  python tests/golden/make_golden_mrstft.py"""
import os
import sys
import warnings

os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
sys.dont_write_bytecode = True
warnings.filterwarnings("ignore")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, "/root/reference/src")

import numpy as np
import torch

from loss import MultiResolutionSTFTLoss            # the reference itself
from oracle import mrstft_oracle as M

CASES = [(1, 2, 8, 20000), (2, 1, 2, 9001), (3, 3, 8, 4500)]


def main():
    out = {}
    ref = MultiResolutionSTFTLoss()
    for seed, B, C, T in CASES:
        x, y = M.pair(seed, B, C, T)
        x.requires_grad_(True)
        loss = ref(x, y)
        loss.backward()
        out[f"loss/{seed}/{B}/{C}/{T}"] = np.array([loss.item()], dtype=np.float64)
        out[f"grad/{seed}/{B}/{C}/{T}/sub7"] = x.grad[..., ::7].numpy()
        out[f"grad/{seed}/{B}/{C}/{T}/absmax"] = np.array([x.grad.abs().max().item()], dtype=np.float64)
    one = MultiResolutionSTFTLoss(fft_sizes=[256], hop_sizes=[64], win_sizes=[256])
    x, y = M.pair(9, 1, 2, 3000)
    out["loss_single256/9/1/2/3000"] = np.array([one(x, y).item()], dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "mrstft_outputs.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
