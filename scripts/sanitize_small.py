"""Small shapes of every kernel family for compute-sanitizer (racecheck / memcheck):
   compute-sanitizer --tool racecheck python scripts/sanitize_small.py
K1 (TMA hand-off through mbarriers + the arrival counter, edge tiles, unaligned rows), K2, K3, K4/K5 (2^19 and
standalone), K1G, K6, K8."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(3)
B, T = 2, 20480 + 77
stems = {n: (torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for n in pkg.STEMS}
fe = pkg.MixingFeatureExtractor()
f, lm = fe.features_and_logmel(stems)                                   # K1 (TMA staging) + K2
wide = torch.rand(B, 8, T + 5, device=dev, generator=g) * 0.1
views = {n: wide[:, 2 * i:2 * i + 2, 3:3 + T] for i, n in enumerate(pkg.STEMS)}
f2, lm2 = fe.features_and_logmel(views)                                 # K1 cooperative staging (unaligned rows)
aug = pkg.AudioAugmenter(prob=1.0)
torch.manual_seed(5)
neg = aug.apply(stems, aug.draw_batch(B, impulse_device=dev))           # K3 + K4/K5 (2^19 transform)
y = aug.apply_reverb(stems["vocals"], 0.1)                              # K4 standalone
z = aug.apply_compression(stems["bass"], -18, 3)                        # K3 single tensor
ev = pkg.MixingFeatureExtractor(44100, 2048, 512, 80)
f3 = ev.extract_all_features(stems)                                     # K1G + K2
enc = pkg.MelSpectrogramPreprocessor()
req = {n: v.clone().requires_grad_(True) for n, v in stems.items()}
enc(req).sum().backward()                                               # K6
loss = pkg.MultiResolutionSTFTLoss()
x = stems["vocals"].clone().requires_grad_(True)
loss(x, stems["drums"]).backward()                                      # K8
torch.cuda.synchronize()
print("ok", float(f.sum()), float(f2.sum()), float(neg["other"].abs().mean()), float(y.abs().mean()), float(z.abs().mean()),
      float(f3.sum()), float(req["bass"].grad.abs().mean()), float(x.grad.abs().mean()))
