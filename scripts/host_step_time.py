"""Host time of one bench step (config 3) against its device time: how far the host runs ahead of the GPU.
Prints the wall time the host needs to QUEUE n steps (no sync inside) and the device time of the same steps."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mixing_style_transfer_b200 as pkg

N_TRI, T = 24, 441000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
stems = [(torch.rand(3 * N_TRI, 2, T, device=dev, generator=g) * 2 - 1) * 0.2 for _ in pkg.STEMS]
fe = pkg.MixingFeatureExtractor(); plan = fe._plan(dev)
aug = pkg.AudioAugmenter(prob=0.5)
anchors = {n: s[:N_TRI] for n, s in zip(pkg.STEMS, stems)}
negatives = {n: s[2 * N_TRI:] for n, s in zip(pkg.STEMS, stems)}
f_out = torch.empty(3 * N_TRI, plan.feature_dim, device=dev)
lm_out = torch.empty(3 * N_TRI, 8, 128, plan.num_frames(T), device=dev)
torch.manual_seed(7)
parts = {"draw": 0.0, "apply": 0.0, "features": 0.0}

def step():
    t0 = time.perf_counter(); dec = aug.draw_batch(N_TRI, impulse_device=dev)
    t1 = time.perf_counter(); aug.apply(anchors, dec, out=negatives)
    t2 = time.perf_counter(); plan.features_logmel(stems, True, True, features_out=f_out, logmel_out=lm_out)
    t3 = time.perf_counter()
    parts["draw"] += t1 - t0; parts["apply"] += t2 - t1; parts["features"] += t3 - t2

for _ in range(5):
    step()
torch.cuda.synchronize()
for n in (20, 20, 20):
    for k in parts: parts[k] = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); t0 = time.perf_counter()
    for _ in range(n):
        step()
    t_host = (time.perf_counter() - t0) / n * 1e3
    e1.record(); torch.cuda.synchronize()
    print(f"host queues a step in {t_host:.3f} ms (draw {parts['draw']/n*1e3:.3f}, apply {parts['apply']/n*1e3:.3f}, "
          f"features {parts['features']/n*1e3:.3f}); device {e0.elapsed_time(e1)/n:.3f} ms per step")
