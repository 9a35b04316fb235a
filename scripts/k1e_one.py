"""A few K1E launches with statistics ((2048, 512, 80) features + log-mel, 72 clips x 10 s) for ncu captures."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg
dev = torch.device("cuda", 0)
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = [(torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for _ in range(4)]
plan = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)._plan(dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    f, lm = plan.features_logmel(stems, True, True)
torch.cuda.synchronize()
print("ok", float(f.sum()))
