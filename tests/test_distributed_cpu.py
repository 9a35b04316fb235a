"""CPU, world_size 2, gloo: the host-side logic of the clip-sharded path (sharding, in-place all-gather)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mixing_style_transfer_b200.distributed import GlobalBatch, shard_bounds, shard_stems


def test_shard_bounds_cover_and_balance():
    for n in (0, 1, 7, 72, 576, 4096):
        for w in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(8, 2, 2)


def test_shard_stems_are_views():
    x = {"vocals": torch.arange(24.).reshape(6, 2, 2)}
    s = shard_stems(x, 1, 3)
    assert s["vocals"].shape == (2, 2, 2) and s["vocals"].data_ptr() == x["vocals"][2:4].data_ptr()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        gb = GlobalBatch(n_local=3, feature_dim=64, n_mels=4, n_frames=5, device="cpu")
        f, lm = gb.local_views()
        f.fill_(float(rank + 1))
        lm.copy_(torch.arange(lm.numel(), dtype=torch.float32).reshape(lm.shape) + 1000 * rank)
        gb.all_gather()
        ok = True
        for r in range(world):
            ok &= bool((gb.features[3 * r:3 * r + 3] == r + 1).all())
            ok &= bool(gb.logmel[3 * r, 0, 0, 0] == 1000 * r)
        ok &= gb.payload_bytes_received() == 4 * 3 * (64 + 8 * 4 * 5) * (world - 1)
        # chunk-major layout: chunk c of every rank is contiguous, gathered chunk by chunk
        gc = GlobalBatch(n_local=6, feature_dim=4, n_mels=2, n_frames=3, device="cpu", n_chunks=3)
        for c in range(3):
            f, lm = gc.local_views(c)
            for i in range(gc.cs):
                f[i].fill_(100.0 * rank + c * gc.cs + i)          # value encodes (rank, local clip)
                lm[i].fill_(100.0 * rank + c * gc.cs + i)
            gc.all_gather(chunk=c)
        for r in range(world):
            for j in range(6):
                row = gc.global_index(r, j)
                ok &= bool((gc.features[row] == 100.0 * r + j).all()) and bool((gc.logmel[row] == 100.0 * r + j).all())
        ok &= sorted(gc.global_index(r, j) for r in range(world) for j in range(6)) == list(range(6 * world))
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_in_place_all_gather_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]
