"""Does an nvidia-smi poller running beside the bench stall its launches?  40 repeats of a 20-step timed region (config 3,
device-timed like bench.py) without a poller, with `-lms 50` and with `-lms 200`; prints the sorted per-repeat ms per step."""
import os, subprocess, sys, time
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


def step():
    aug.apply(anchors, aug.draw_batch(N_TRI, impulse_device=dev), out=negatives)
    plan.features_logmel(stems, True, True, features_out=f_out, logmel_out=lm_out)


def region(n=20):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
     "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
for _ in range(10):
    step()
for label, lms in (("no poller", None), ("-lms 50", 50), ("-lms 200", 200), ("no poller", None)):
    proc = None
    if lms:
        proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={Q}", "--format=csv,noheader,nounits", "-i", "0", "-lms", str(lms)],
                                stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        time.sleep(3.0)
    xs = sorted(region() for _ in range(40))
    if proc:
        proc.terminate(); proc.wait()
    print(f"{label:10s}: min {xs[0]:.3f}  median {xs[20]:.3f}  p90 {xs[36]:.3f}  max {xs[-1]:.3f}   worst five: " + " ".join(f"{x:.2f}" for x in xs[-5:]))
