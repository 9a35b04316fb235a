"""Clip-sharded data parallelism for the DSP front end: one process per GPU (torchrun), a contiguous
block of clips per rank, and ONE collective -- the all-gather of per-clip features and log-mels into
the global contrastive batch.  The reference has no distributed code at all
(single process, `torch.device(args.device)`, src/train.py:439); this is the new-framework equivalent
BASELINE.json asks for.

Every output of the path depends on exactly one clip, so there is no data-path collective: the kernels
write their results straight into this rank's slice of the pre-allocated global buffers and
`all_gather_into_tensor` runs in place over NCCL / NVLink (gloo on CPU tensors in the tests).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_clips: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of clips owned by `rank`; the first n % world ranks get one extra."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(n_clips, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_stems(stems_dict: Dict[str, torch.Tensor], rank: int, world_size: int) -> Dict[str, torch.Tensor]:
    """Views (no copy) of this rank's clips of a dict of (N,2,T) tensors.  A triplet's anchor/positive/
    negative stay together when the caller orders clips triplet-major and N % (3*world) == 0."""
    n = next(iter(stems_dict.values())).shape[0]
    lo, hi = shard_bounds(n, rank, world_size)
    return {k: v[lo:hi] for k, v in stems_dict.items()}


class GlobalBatch:
    """Pre-allocated global buffers (N_global, D) and (N_global, 8, n_mels, F) on this rank's device.
    `local_views()` are the slices the kernels write; `all_gather()` fills the rest in place."""

    def __init__(self, n_local: int, feature_dim: int, n_mels: int, n_frames: int, device,
                 group: Optional[dist.ProcessGroup] = None, with_logmel: bool = True):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n_local = n_local
        n = n_local * self.world                    # all_gather_into_tensor needs equal shards
        self.features = torch.empty((n, feature_dim), dtype=torch.float32, device=device)
        self.logmel = torch.empty((n, 8, n_mels, n_frames), dtype=torch.float32, device=device) if with_logmel else None

    def local_views(self):
        lo = self.rank * self.n_local
        f = self.features[lo:lo + self.n_local]
        lm = self.logmel[lo:lo + self.n_local] if self.logmel is not None else None
        return f, lm

    def payload_bytes_received(self) -> int:
        per = self.features[0].numel() + (self.logmel[0].numel() if self.logmel is not None else 0)
        return 4 * per * self.n_local * (self.world - 1)

    def all_gather(self, async_op: bool = False):
        """In-place all-gather of both buffers (input = this rank's slice of the output)."""
        if self.world == 1:
            return []
        f, lm = self.local_views()
        works = [dist.all_gather_into_tensor(self.features, f, group=self.group, async_op=async_op)]
        if lm is not None:
            works.append(dist.all_gather_into_tensor(self.logmel, lm, group=self.group, async_op=async_op))
        return works


def extract_global_batch(extractor, local_stems: Dict[str, torch.Tensor], group=None, with_logmel: bool = True,
                         gather: bool = True) -> GlobalBatch:
    """features + log-mel of this rank's clips written into the global buffers, then all-gathered."""
    from ._native import STEMS
    first = local_stems[STEMS[0]]
    n_local, _, T = first.shape
    plan = extractor._plan(first.device)
    gb = GlobalBatch(n_local, plan.feature_dim, plan.n_mels, plan.num_frames(T), first.device, group, with_logmel)
    f, lm = gb.local_views()
    plan.features_logmel([local_stems[n] for n in STEMS], True, with_logmel, features_out=f, logmel_out=lm)
    if gather:
        gb.all_gather()
    return gb
