"""Copy-engine timeline of the end-to-end loop (TripletPipeline.submit/wait, config 3) from CUPTI: busy time, bytes and
rate of each direction, the idle gaps of the upload stream, and the kernels' share.  Usage: e2e_timeline.py [chunk] [out]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
import mixing_style_transfer_b200 as pkg
from mixing_style_transfer_b200.pipeline import HostPipeline, TripletPipeline

N_TRI, T = 24, 441000
chunk = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = torch.device("cuda:0")
fe = pkg.MixingFeatureExtractor()
plan = fe._plan(dev)
F = plan.num_frames(T)
if os.environ.get("E2E_ONE_ALLOC", "1") != "0":
    host_all = HostPipeline.pinned_like((4, 2 * N_TRI, 2, T))
    host = {n: host_all[k] for k, n in enumerate(pkg.STEMS)}
else:
    host = {n: HostPipeline.pinned_like((2 * N_TRI, 2, T)) for n in pkg.STEMS}
for n in pkg.STEMS:
    host[n].uniform_(-0.2, 0.2)
hfs = [HostPipeline.pinned_like((3 * N_TRI, 64)) for _ in range(2)]
hls = [HostPipeline.pinned_like((3 * N_TRI, 8, 128, F)) for _ in range(2)]
pipe = TripletPipeline(fe, pkg.AudioAugmenter(prob=0.5), N_TRI, T, chunk=chunk, device=dev, with_logmel=True)


def loop(n):
    pending = pipe.submit(host, pipe.aug.draw_batch(N_TRI, impulse_device=dev), hfs[0], hls[0])
    for k in range(1, n):
        nxt = pipe.submit(host, pipe.aug.draw_batch(N_TRI, impulse_device=dev), hfs[k & 1], hls[k & 1])
        pipe.wait(pending)
        pending = nxt
    pipe.wait(pending)


loop(3)
torch.cuda.synchronize()
N = 4
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    loop(N)
    torch.cuda.synchronize()
ev = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e.time_range.start)
t0, t1 = ev[0].time_range.start, max(e.time_range.end for e in ev)
lines = [f"chunk {chunk} triplets, {N} steps: span {(t1 - t0) / N / 1e3:.3f} ms per step"]
for tag, nbytes in (("Memcpy HtoD", pipe.h2d_bytes), ("Memcpy DtoH", pipe.d2h_bytes)):
    xs = [e for e in ev if e.name.startswith(tag) and (e.time_range.end - e.time_range.start) > 20]
    busy = sum(e.time_range.end - e.time_range.start for e in xs)
    gaps = sorted(((b.time_range.start - a.time_range.end, a.time_range.end - t0) for a, b in zip(xs, xs[1:])), reverse=True)
    lines.append(f"{tag}: {len(xs)} copies, busy {busy / N / 1e3:.3f} ms per step -> {nbytes * N / busy / 1e3:.1f} GB/s while busy; "
                 f"idle {(t1 - t0 - busy) / N / 1e3:.3f} ms per step; largest gaps (us @ us): "
                 + ", ".join(f"{g:.0f}@{at:.0f}" for g, at in gaps[:6]))
ks = [e for e in ev if "mixfe::" in e.name]
lines.append(f"kernels: {sum(e.time_range.end - e.time_range.start for e in ks) / N / 1e3:.3f} ms per step in {len(ks) // N} launches")
out = "\n".join(lines)
print(out)
if len(sys.argv) > 2:
    open(sys.argv[2], "a").write(out + "\n")
if os.environ.get("E2E_DUMP"):
    with open(os.environ["E2E_DUMP"], "w") as f:
        for e in ev:
            d = e.time_range.end - e.time_range.start
            if e.time_range.start - t0 > 2 * (t1 - t0) / N:
                break
            f.write(f"{e.time_range.start - t0:10.1f} {d:9.1f} {e.name[:60]}\n")
