"""A handful of K1 launches on the bench batch (for ncu captures): python scripts/k1_one.py [n_launches]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg
dev = torch.device("cuda", 0)
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = [(torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for _ in range(4)]
plan = pkg.MixingFeatureExtractor()._plan(dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5
f = lm = None
for _ in range(n):
    f, lm = plan.features_logmel(stems, True, True, features_out=f, logmel_out=lm)
torch.cuda.synchronize()
print("ok", float(f.sum()), float(lm.mean()))
