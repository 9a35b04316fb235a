"""A few augmentation calls on the bench's shape (for ncu captures): python scripts/aug_one.py [n_calls] [prob]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg
dev = torch.device("cuda", 0)
B, T = 24, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = {n: (torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for n in pkg.STEMS}
outs = {n: torch.empty_like(v) for n, v in stems.items()}
aug = pkg.AudioAugmenter(prob=float(sys.argv[2]) if len(sys.argv) > 2 else 0.5)
torch.manual_seed(1234)
dec = aug.draw_batch(B, impulse_device=dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    aug.apply(stems, dec, out=outs)
torch.cuda.synchronize()
print("ok", float(outs["vocals"].abs().mean()))
