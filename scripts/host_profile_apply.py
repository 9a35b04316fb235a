"""cProfile of the host side of AudioAugmenter.draw_batch + apply (24 clips): where the 1.2 ms of host time per step go."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mixing_style_transfer_b200 as pkg

N_TRI, T = 24, 441000
dev = torch.device("cuda:0")
stems = [torch.rand(2 * N_TRI, 2, T, device=dev) * 0.2 for _ in pkg.STEMS]
aug = pkg.AudioAugmenter(prob=0.5)
anchors = {n: s[:N_TRI] for n, s in zip(pkg.STEMS, stems)}
negatives = {n: s[N_TRI:] for n, s in zip(pkg.STEMS, stems)}
for _ in range(5):
    aug.apply(anchors, aug.draw_batch(N_TRI, impulse_device=dev), out=negatives)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(100):
    aug.apply(anchors, aug.draw_batch(N_TRI, impulse_device=dev), out=negatives)
    if _ % 4 == 3:
        torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
