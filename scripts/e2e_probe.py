"""Where the end-to-end time goes: raw pinned H2D / D2H rates alone and together, then HostPipeline at several chunk sizes."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg
from mixing_style_transfer_b200.pipeline import HostPipeline
dev = torch.device("cuda", 0)
N, T = 72, 441000
h = torch.empty((N, 8, T), dtype=torch.float32, pin_memory=True).normal_()
d = torch.empty_like(h, device=dev)
ho = torch.empty((N, 8 * 128 * 1723), dtype=torch.float32, pin_memory=True)
do = torch.empty_like(ho, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
def up():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
def down():
    with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
def both(): up(); down()
a = t(up); b = t(down); c = t(both)
print(f"H2D 1.016 GB alone {a:.2f} ms ({h.numel()*4/a/1e6:.1f} GB/s); D2H 0.508 GB alone {b:.2f} ms ({ho.numel()*4/b/1e6:.1f} GB/s); together {c:.2f} ms")
fe = pkg.MixingFeatureExtractor()
host = {n: HostPipeline.pinned_like((N, 2, T)).normal_(0, 0.1) for n in pkg.STEMS}
hf = HostPipeline.pinned_like((N, 64)); hl = HostPipeline.pinned_like((N, 8, 128, 1723))
for chunk in (3, 6, 12, 24):
    pipe = HostPipeline(fe, N, T, chunk=chunk, device=dev)
    print(f"HostPipeline chunk={chunk}: {t(lambda: pipe.run(host, hf, hl), 4):.2f} ms")
    del pipe
