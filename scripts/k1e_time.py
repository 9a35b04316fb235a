import os, sys, torch
sys.path.insert(0, "/root/repo")
import mixing_style_transfer_b200 as pkg
dev = torch.device("cuda", 0)
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = [(torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for _ in range(4)]
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
plan = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)._plan(dev)
print(plan.k1_config())
for _ in range(2):
    print("logmel only", timed(lambda: plan.features_logmel(stems, False, True)))
    print("features+logmel", timed(lambda: plan.features_logmel(stems, True, True)))
    print("features only", timed(lambda: plan.features_logmel(stems, True, False)))
