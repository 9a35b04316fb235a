"""Time K1 in isolation on the bench batch for the four (features, logmel) output combinations."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg
dev = torch.device("cuda", 0)
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = [(torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for _ in range(4)]
fe = pkg.MixingFeatureExtractor()
plan = fe._plan(dev)
print(plan.k1_config())
for wf, wl in ((True, True), (False, True), (True, False)):
    for _ in range(3):
        plan.features_logmel(stems, wf, wl)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        plan.features_logmel(stems, wf, wl)
    e1.record(); torch.cuda.synchronize()
    print(f"features={wf} logmel={wl}: {e0.elapsed_time(e1)/20:.3f} ms")
