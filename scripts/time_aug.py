"""Time the augmentation kernels on the bench's shape: the 24 anchors of the 72-clip batch -> 24 negatives
(prob 0.5 decisions from draw_batch, seed 1234).  CUDA events inside libmixfe (K3, K4/K5) and around the call.
MIXFE_LIB=<other build> selects a kernel variant."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg

dev = torch.device("cuda", 0)
B, T = 24, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = {n: (torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for n in pkg.STEMS}
outs = {n: torch.empty_like(v) for n, v in stems.items()}
for prob in (0.5, 1.0):
    aug = pkg.AudioAugmenter(prob=prob)
    torch.manual_seed(1234)
    dec = aug.draw_batch(B, impulse_device=dev)
    plan = aug._plan(dev)
    for _ in range(3):
        aug.apply(stems, dec, out=outs)
    torch.cuda.synchronize()
    plan.enable_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        aug.apply(stems, dec, out=outs)
    e1.record(); torch.cuda.synchronize()
    k3, n3 = plan.kernel_time("augment")
    k4, _ = plan.kernel_time("reverb")
    plan.enable_timing(False)
    byts = 2 * B * 8 * T * 4
    print(f"lib={os.environ.get('MIXFE_LIB', 'default')} prob={prob}: K3 {k3 / n3:.4f} ms ({byts / (k3 / n3) / 1e6:.0f} GB/s), "
          f"K4/K5 {k4 / n:.4f} ms per step ({len(dec.reverb_rows)} reverbs), call {e0.elapsed_time(e1) / n:.4f} ms")
