"""First-light check of K1/K2 on a GPU box: mel power, log-mel and features vs the oracle."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mixing_style_transfer_b200 as pkg
from oracle import mixfe_oracle as O

dev = torch.device("cuda:0")
print(torch.cuda.get_device_name(0))
for mk, T in ((O.synthetic_clip, 4096), (O.musical_clip, 44100), (O.synthetic_clip, 66150)):
    clip = mk(1234, T)
    fe = pkg.MixingFeatureExtractor()
    pl = fe._plan(dev)
    st = [clip[n][None].to(dev) for n in pkg.STEMS]
    mp = pl.debug_mel_power(st).cpu()
    torch.cuda.synchronize()
    ofe = O.MelFrontEnd(O.FrontEndConfig())
    ref = torch.cat([ofe.mel_power(clip[n][None]) for n in pkg.STEMS], dim=1)
    err = (mp - ref).abs()
    fmax = ref.amax(dim=2, keepdim=True)
    rel = (err / (1e-4 * ref.abs() + 1e-6 * fmax + 1e-30)).max()
    print(mk.__name__, T, "melpow max abs", float(err.max()), "ref max", float(ref.max()), "tol-units", float(rel))
    if float(rel) > 1:
        idx = torch.nonzero(err / (1e-4 * ref.abs() + 1e-6 * fmax + 1e-30) > 1)
        print(" bad count", idx.shape[0], "first", idx[:10].tolist())
        for i in idx[:5].tolist():
            print("  ", i, float(mp[tuple(i)]), float(ref[tuple(i)]))
    feats, lm = fe.features_and_logmel({n: clip[n].to(dev) for n in pkg.STEMS})
    rf = O.mixing_features(clip)
    rl = O.log_mel({n: clip[n][None] for n in pkg.STEMS})
    print("  logmel max abs err", float((lm.cpu() - rl).abs().max()))
    d = (feats.cpu() - rf).abs()
    tol = 1e-4 * rf.abs() + 1e-5
    names = fe.feature_names()
    bad = torch.nonzero(d > tol).flatten().tolist()
    print("  features max abs err", float(d.max()), "bad", len(bad))
    for i in bad[:20]:
        print("    ", i, names[i], float(feats[i]), float(rf[i]))

# quick timing at the bench shape
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
st = [(torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for _ in range(4)]
fe = pkg.MixingFeatureExtractor()
pl = fe._plan(dev)
for _ in range(3):
    f, l = pl.features_logmel(st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    f, l = pl.features_logmel(st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"72 clips x 10 s: {ms:.3f} ms/step -> {B*10/ms*1e3:.0f} audio-s/s; alg bytes {72*21169664/ms/1e6:.0f} GB/s")
