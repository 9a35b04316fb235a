"""Time the multi-resolution STFT loss (K8), forward + backward, on a training-sized batch (16 x 2 x 10 s), next to
the same loss written with torch.stft / autograd on the same GPU (the code path of the reference's module,
src/loss.py:332-448, restated in plain torch here because /root/reference is not on the GPU box).
CUDA events, 3 warm-up + 10 timed iterations."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg

dev = torch.device("cuda", 0)
B, C, T = 16, 2, 441000
g = torch.Generator(device=dev).manual_seed(3)
x = ((torch.rand(B, C, T, device=dev, generator=g) * 2 - 1) * 0.25).requires_grad_(True)
y = (torch.rand(B, C, T, device=dev, generator=g) * 2 - 1) * 0.25
FFT, HOP = (1024, 2048, 512), (256, 512, 128)


def torch_loss(x, y):
    total = 0.0
    for n, h in zip(FFT, HOP):
        w = torch.hann_window(n, device=x.device)
        X = torch.stft(x.reshape(-1, x.shape[-1]), n, h, n, w, return_complex=True).abs()
        Y = torch.stft(y.reshape(-1, y.shape[-1]), n, h, n, w, return_complex=True).abs()
        sc = torch.norm(Y - X, p="fro") / (torch.norm(Y, p="fro") + 1e-8)
        lm = torch.nn.functional.l1_loss(torch.log(X + 1e-5), torch.log(Y + 1e-5))
        total = total + sc + lm
    return total / len(FFT)


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


mod = pkg.MultiResolutionSTFTLoss(FFT, HOP, FFT)


def step(fn):
    x.grad = None
    fn(x, y).backward()


a, b = float(mod(x, y).detach()), float(torch_loss(x, y).detach())
print(f"loss: K8 {a:.6f}  torch {b:.6f}")
print(f"K8    forward+backward: {timed(lambda: step(mod)):.3f} ms")
print(f"K8    forward only:     {timed(lambda: mod(x.detach(), y)):.3f} ms")
print(f"torch forward+backward: {timed(lambda: step(torch_loss)):.3f} ms   (peak memory {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB)")

# per-resolution kernel times (CUDA events inside libmixfe around the loss kernels of each plan)
plans = mod._plans_for(dev)
for pl in plans:
    pl.enable_timing(True)
for _ in range(10):
    step(mod)
torch.cuda.synchronize()
for (n, h), pl in zip(zip(FFT, HOP), plans):
    ms, cnt = pl.kernel_time("stft_loss")
    pl.enable_timing(False)
    print(f"  resolution ({n}, {h}): {ms / 10:.3f} ms per forward + backward ({cnt} timed brackets)")
