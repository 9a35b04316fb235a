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
    `local_views()` are the slices the kernels write; `all_gather()` fills the rest in place.

    Layout.  With n_chunks == 1 the global order is rank-major (rank r owns rows [r*n_local, (r+1)*n_local)).
    With n_chunks > 1 it is chunk-major: chunk c of every rank sits together,
        row = c * (world * cs) + r * cs + i,   cs = n_local / n_chunks,
    so that chunk c can be all-gathered in place (equal contiguous shards) as soon as it has been computed,
    while the kernels work on chunk c + 1 (`extract_global_batch(..., n_chunks=k)`).  `global_index()` maps a
    (rank, local clip) pair to its row either way."""

    def __init__(self, n_local: int, feature_dim: int, n_mels: int, n_frames: int, device,
                 group: Optional[dist.ProcessGroup] = None, with_logmel: bool = True, n_chunks: int = 1,
                 symmetric=False):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if n_chunks < 1 or n_local % n_chunks:
            raise ValueError(f"n_chunks={n_chunks} must divide n_local={n_local}")
        self.n_local = n_local
        self.n_chunks = n_chunks
        self.cs = n_local // n_chunks
        n = n_local * self.world                    # all_gather_into_tensor needs equal shards
        self._peers = None                          # symmetric mode: (features, logmel) views of every rank's buffers
        self._handle = None
        self._mc = None                             # multicast mode: (address of features, address of logmel) in the multicast mapping
        self._stage = None                          # multicast mode: this rank's results before they are multicast
        self.push_sms = 0                           # symmetric="sm": CTAs of the SM-driven push kernel (0: copy engines)
        if symmetric and self.world > 1:
            # One NVLink-mapped allocation per rank (torch symmetric memory); the gather is then a set of
            # copy-engine pushes into the peers' buffers: no SMs, so it runs under the persistent feature kernel.
            import torch.distributed._symmetric_memory as symm
            nf, nl = n * feature_dim, (n * 8 * n_mels * n_frames if with_logmel else 0)
            flat = symm.empty((nf + nl,), dtype=torch.float32, device=device)
            grp = group if group is not None else dist.group.WORLD
            self._handle = symm.rendezvous(flat, grp.group_name)

            def views(t):
                f = t[:nf].view(n, feature_dim)
                return f, (t[nf:].view(n, 8, n_mels, n_frames) if with_logmel else None)
            self.features, self.logmel = views(flat)
            self._peers = [views(self._handle.get_buffer(r, (nf + nl,), torch.float32)) for r in range(self.world)]
            if symmetric == "sm":
                self.push_sms = 16
            if symmetric == "multicast":
                # One copy to the NVSwitch multicast mapping lands in EVERY rank's buffer (this one included), so the
                # kernels write a local staging slice and 2 copies per chunk replace the 2 (world - 1) pushes.
                if not self._handle.multicast_ptr:
                    raise RuntimeError("symmetric='multicast': this allocation has no NVSwitch multicast mapping")
                self._mc = (self._handle.multicast_ptr, self._handle.multicast_ptr + 4 * nf)
                self._stage = (torch.empty((n_local, feature_dim), dtype=torch.float32, device=device),
                               torch.empty((n_local, 8, n_mels, n_frames), dtype=torch.float32, device=device) if with_logmel else None)
            return
        self.features = torch.empty((n, feature_dim), dtype=torch.float32, device=device)
        self.logmel = torch.empty((n, 8, n_mels, n_frames), dtype=torch.float32, device=device) if with_logmel else None

    def global_index(self, rank: int, local_clip: int) -> int:
        c, i = divmod(local_clip, self.cs)
        return c * self.world * self.cs + rank * self.cs + i

    def _chunk_span(self, c: int) -> Tuple[int, int]:
        lo = c * self.world * self.cs
        return lo, lo + self.world * self.cs

    def local_views(self, chunk: Optional[int] = None):
        """This rank's slice of the global buffers: of chunk `chunk`, or (n_chunks == 1 only) of everything."""
        if chunk is None:
            if self.n_chunks != 1:
                raise ValueError("chunk-major layout: ask for one chunk's views")
            chunk = 0
        if self._stage is not None:
            a = chunk * self.cs
            return self._stage[0][a:a + self.cs], (self._stage[1][a:a + self.cs] if self._stage[1] is not None else None)
        lo = self._chunk_span(chunk)[0] + self.rank * self.cs
        f = self.features[lo:lo + self.cs]
        lm = self.logmel[lo:lo + self.cs] if self.logmel is not None else None
        return f, lm

    def payload_bytes_received(self) -> int:
        per = self.features[0].numel() + (self.logmel[0].numel() if self.logmel is not None else 0)
        return 4 * per * self.n_local * (self.world - 1)

    def push(self, chunk: Optional[int] = None) -> None:
        """Symmetric mode: copy this rank's slice (of one chunk, or of all) into every peer's buffers on the current
        stream (cudaMemcpyAsync over NVLink peer mappings: copy engines, no kernels).  Call `fence()` once the last
        push is enqueued."""
        for c in (range(self.n_chunks) if chunk is None else [chunk]):
            lo = self._chunk_span(c)[0] + self.rank * self.cs
            f, lm = self.local_views(c)
            if self._mc is not None:
                from ._native import copy_async
                if lm is not None:
                    copy_async(self._mc[1] + 4 * lo * self.logmel[0].numel(), lm, 4 * lm.numel())
                copy_async(self._mc[0] + 4 * lo * self.features.shape[1], f, 4 * f.numel())
                continue
            order = [(self.rank + k) % self.world for k in range(1, self.world)]
            if self.push_sms:
                # one kernel on a few SMs stores the slice into all peers (K11); the feature kernel leaves it the SMs
                from ._native import push_peers
                if lm is not None:
                    push_peers([self._peers[r][1][lo:lo + self.cs].data_ptr() for r in order], lm, 4 * lm.numel(), self.push_sms)
                push_peers([self._peers[r][0][lo:lo + self.cs].data_ptr() for r in order], f, 4 * f.numel(), 1)
                continue
            for r in order:
                pf, plm = self._peers[r]
                if lm is not None:
                    plm[lo:lo + self.cs].copy_(lm, non_blocking=True)
                pf[lo:lo + self.cs].copy_(f, non_blocking=True)

    def fence(self) -> None:
        """Symmetric mode: stream-ordered barrier across ranks -- when it completes on this rank's stream, every push the
        peers enqueued before THEIR fence has landed here."""
        self._handle.barrier(channel=0)

    def all_gather(self, async_op: bool = False, chunk: Optional[int] = None):
        """In-place all-gather of both buffers (input = this rank's slice of the output); one chunk or all."""
        if self.world == 1:
            return []
        if self._peers is not None:
            self.push(chunk)
            if chunk is None:
                self.fence()
            return []
        works = []
        for c in (range(self.n_chunks) if chunk is None else [chunk]):
            lo, hi = self._chunk_span(c)
            f, lm = self.local_views(c)
            works.append(dist.all_gather_into_tensor(self.features[lo:hi], f, group=self.group, async_op=async_op))
            if lm is not None:
                works.append(dist.all_gather_into_tensor(self.logmel[lo:hi], lm, group=self.group, async_op=async_op))
        return works


def extract_global_batch(extractor, local_stems: Dict[str, torch.Tensor], group=None, with_logmel: bool = True,
                         gather: bool = True, n_chunks: int = 1, comm_stream=None,
                         out: Optional[GlobalBatch] = None, symmetric=False) -> GlobalBatch:
    """features + log-mel of this rank's clips written into the global buffers, then all-gathered.

    n_chunks > 1 pipelines the two: the clips are processed in n_chunks launches and the gather of chunk c runs on
    `comm_stream` (a side stream, created when None) while the kernels compute chunk c + 1; the current stream waits
    for the last gather before returning.  The global order is then chunk-major (GlobalBatch.global_index).
    With NCCL the overlap is small -- its kernels need whole SMs and the persistent feature kernel holds every
    register of every SM -- so `symmetric=True` (copy-engine pushes into NVLink-mapped peer buffers, GlobalBatch.push)
    is the mode that pipelines; `symmetric="multicast"` sends each chunk ONCE, to the NVSwitch multicast mapping.
    `out` reuses a GlobalBatch (symmetric buffers cost a rendezvous to create); the
    caller then makes sure the peers are done reading the previous contents."""
    from ._native import STEMS
    first = local_stems[STEMS[0]]
    n_local, _, T = first.shape
    plan = extractor._plan(first.device)
    gb = out if out is not None else GlobalBatch(n_local, plan.feature_dim, plan.n_mels, plan.num_frames(T), first.device,
                                                 group, with_logmel, n_chunks=n_chunks, symmetric=symmetric)
    if gb.n_local != n_local or gb.n_chunks != n_chunks:
        raise ValueError("`out` was built for another batch shape or chunk count")
    side = gather and gb.world > 1 and first.is_cuda and n_chunks > 1
    if side and comm_stream is None:
        comm_stream = torch.cuda.Stream(device=first.device)
    reserve = gb.push_sms if (side and gb.push_sms) else 0
    if reserve:
        plan.reserve_sms(reserve)                   # K1's persistent grid leaves the push kernel its SMs
    try:
        return _extract_chunks(plan, gb, local_stems, STEMS, n_chunks, with_logmel, gather, side, comm_stream, first)
    finally:
        if reserve:
            plan.reserve_sms(0)


def _extract_chunks(plan, gb, local_stems, STEMS, n_chunks, with_logmel, gather, side, comm_stream, first):
    for c in range(n_chunks):
        f, lm = gb.local_views(c)
        sl = slice(c * gb.cs, (c + 1) * gb.cs)
        plan.features_logmel([local_stems[n][sl] for n in STEMS], True, with_logmel, features_out=f, logmel_out=lm)
        if side:
            done = torch.cuda.Event()
            done.record()
            with torch.cuda.stream(comm_stream):
                comm_stream.wait_event(done)
                gb.all_gather(chunk=c)
        elif gather:
            gb.all_gather(chunk=c)
    if gather and gb._peers is not None:
        if side:
            with torch.cuda.stream(comm_stream):
                gb.fence()
        else:
            gb.fence()
    if side:
        torch.cuda.current_stream(first.device).wait_stream(comm_stream)
    return gb
