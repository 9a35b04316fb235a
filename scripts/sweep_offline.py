"""BASELINE.json config 4: offline preprocessing sweep -- N synthetic 30 s 8-channel clips, features + log-mel,
sharded contiguously by clip index over the ranks (no collective, outputs stay sharded).

    python scripts/sweep_offline.py --clips 4096                              # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/sweep_offline.py --clips 4096

Clip i is regenerated on the device from seed = i (so a clip's audio, and hence its features, do not depend on the
world size: `checksum` -- the fp64 sum over all clips of each clip's feature vector, all-reduced -- is the same at
every N).  Each rank walks its shard in resident batches of `--batch` clips; the timed region (CUDA events, max over
ranks) is the extraction kernels of every batch, the regeneration of the next batch sits outside it and its wall
time is reported next to it.  Log-mels land in one resident (batch, 8, n_mels, F) buffer that the next batch
overwrites (a real sweep writes it to storage there); features are kept for the whole shard."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=4096)
    ap.add_argument("--seconds", type=float, default=30.0)
    ap.add_argument("--batch", type=int, default=24)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import mixing_style_transfer_b200 as pkg
    from mixing_style_transfer_b200.distributed import shard_bounds

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("sweep_offline.py needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    SR, T = 44100, int(round(a.seconds * 44100))
    fe = pkg.MixingFeatureExtractor()
    plan = fe._plan(dev)
    F = 1 + T // fe.hop_length
    lo, hi = shard_bounds(a.clips, rank, world)
    stems = [torch.empty(a.batch, 2, T, device=dev) for _ in pkg.STEMS]
    feats = torch.zeros(hi - lo, plan.feature_dim, device=dev)
    lm = torch.empty(a.batch, 8, fe.n_mels, F, device=dev)
    g = torch.Generator(device=dev)

    def fill(first, n):
        for j in range(n):
            g.manual_seed(first + j)
            for s in stems:
                gain = 10 ** (-12.0 * torch.rand(1, device=dev, generator=g) / 20)
                torch.rand(2, T, device=dev, generator=g, out=s[j])
                s[j].mul_(0.5).sub_(0.25).mul_(gain)

    # warm-up on the first batch (also what the first timed batch processes)
    n = min(a.batch, hi - lo)
    fill(lo, n)
    for _ in range(3):
        plan.features_logmel([s[:n] for s in stems], True, True, features_out=feats[:n], logmel_out=lm[:n])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    n0 = pkg.launch_count()
    events, t_wall = [], time.perf_counter()
    for first in range(lo, hi, a.batch):
        n = min(a.batch, hi - first)
        if first != lo:
            fill(first, n)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.features_logmel([s[:n] for s in stems], True, True, features_out=feats[first - lo:first - lo + n],
                             logmel_out=lm[:n])
        e1.record()
        events.append((e0, e1))
    torch.cuda.synchronize()
    wall = time.perf_counter() - t_wall
    launches = pkg.launch_count() - n0
    ms = sum(e0.elapsed_time(e1) for e0, e1 in events)
    t = torch.tensor([ms, wall * 1e3], device=dev, dtype=torch.float64)
    chk = feats.double().sum(0)
    finite = torch.isfinite(feats).all().to(torch.float64).reshape(1)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(chk)
        dist.all_reduce(finite, op=dist.ReduceOp.MIN)
    if rank == 0:
        ms, wall_ms = float(t[0]), float(t[1])
        audio_s = a.clips * a.seconds
        alg = a.clips * (8 * T * 4 + 8 * fe.n_mels * F * 4 + plan.feature_dim * 4)
        print(json.dumps({
            "metric": "audio-sec/sec features+mel (offline sweep, BASELINE config 4)", "n_gpus": world,
            "clips": a.clips, "clip_seconds": a.seconds, "batch": a.batch, "frames": F,
            "kernel_ms_max_over_ranks": ms, "value": audio_s / (ms * 1e-3), "unit": "audio-s/s",
            "algorithmic_bytes": alg, "achieved_gbs_per_gpu": alg / world / (ms * 1e-3) / 1e9,
            "wall_ms_with_regeneration": wall_ms, "wall_audio_s_per_s": audio_s / (wall_ms * 1e-3),
            "gpu_launches_rank0": launches, "all_finite": bool(finite.item() == 1.0),
            "checksum_head": [float(v) for v in chk[:6]], "checksum_sum": float(chk.sum()),
            "scaling": "strong (fixed 4096-clip job; contiguous shards, no collective on the data path)"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
