"""Device timeline of one bench step (config 3) from CUPTI via torch.profiler: every kernel / memcpy / memset with its
start offset, duration and the idle gap before it.  nsys is not in the image; CUPTI sees our launches too because the
activity API is process-wide.  Usage: python scripts/step_timeline.py [out.txt]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch.profiler import ProfilerActivity, profile

import mixing_style_transfer_b200 as pkg

N_TRI, T = 24, 441000


def main():
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    stems = [(torch.rand(3 * N_TRI, 2, T, device=dev, generator=g) * 2 - 1) * 0.2 for _ in pkg.STEMS]
    fe = pkg.MixingFeatureExtractor()
    plan = fe._plan(dev)
    aug = pkg.AudioAugmenter(prob=0.5)
    anchors = {n: s[:N_TRI] for n, s in zip(pkg.STEMS, stems)}
    negatives = {n: s[2 * N_TRI:] for n, s in zip(pkg.STEMS, stems)}
    f_out = torch.empty(3 * N_TRI, plan.feature_dim, device=dev)
    lm_out = torch.empty(3 * N_TRI, 8, 128, plan.num_frames(T), device=dev)
    torch.manual_seed(7)

    def step():
        dec = aug.draw_batch(N_TRI, impulse_device=dev)
        aug.apply(anchors, dec, out=negatives)
        plan.features_logmel(stems, True, True, features_out=f_out, logmel_out=lm_out)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(6):
            step()
        torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ev.sort(key=lambda e: e.time_range.start)
    # one step = from a k3_augment start to the next
    starts = [i for i, e in enumerate(ev) if "k3_augment" in e.name]
    lines = []
    if len(starts) >= 4:
        lo, hi = starts[2], starts[3]
        t0 = ev[lo].time_range.start
        prev_end = None
        # the step really begins with the impulse draw before K3: walk back over torch kernels
        while lo > 0 and "k2_" not in ev[lo - 1].name and "k1_" not in ev[lo - 1].name:
            lo -= 1
        t0 = ev[lo].time_range.start
        busy = 0.0
        for e in ev[lo:hi]:
            if "k3_augment" in e.name and e is not ev[starts[2]]:
                break
            s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
            gap = 0.0 if prev_end is None else e.time_range.start - prev_end
            prev_end = max(prev_end or 0, e.time_range.end)
            busy += d
            lines.append(f"{s:9.1f} us  dur {d:8.1f}  gap {gap:6.1f}  {e.name[:90]}")
        # step period = distance between consecutive first events
        nxt = hi
        while nxt > 0 and "k2_" not in ev[nxt - 1].name and "k1_" not in ev[nxt - 1].name:
            nxt -= 1
        period = ev[nxt].time_range.start - t0
        lines.append(f"step period {period:.1f} us, device busy {busy:.1f} us, idle {period - busy:.1f} us")
    # host side: duration of the python-visible calls
    out = "\n".join(lines)
    print(out)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            f.write(out + "\n")


if __name__ == "__main__":
    main()
