"""Time the secondary kernels on the bench batch (72 clips x 10 s): K1G on the reference's evaluation configuration
(2048, 512, 80) and K6 (log-mel backward) on both configurations.  CUDA events, 3 warm-up + 10 timed launches."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg

dev = torch.device("cuda", 0)
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = [(torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for _ in range(4)]


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for cfg in ((1024, 256, 128), (2048, 512, 80)):
    fe = pkg.MixingFeatureExtractor(44100, *cfg)
    plan = fe._plan(dev)
    F = plan.num_frames(T)
    print(cfg, plan.k1_config())
    print(f"  features+logmel: {timed(lambda: plan.features_logmel(stems, True, True)):.3f} ms")
    print(f"  logmel only:     {timed(lambda: plan.features_logmel(stems, False, True)):.3f} ms")
    gl = torch.randn(B, 8, cfg[2], F, device=dev, generator=g)
    _, lm = plan.features_logmel(stems, False, True)
    for label, kw in (("generic K6 (no forward output)", {}), ("with the forward's log-mel (K6F at 1024/256, K6E at 2048/512)", {"logmel": lm})):
        plan.enable_timing(True)
        total = timed(lambda: plan.logmel_backward(stems, gl, **kw))
        ms, n = plan.kernel_time("logmel_bwd")
        plan.enable_timing(False)
        print(f"  logmel backward, {label}: kernels {ms / n:.3f} ms, call incl. allocation {total:.3f} ms")
