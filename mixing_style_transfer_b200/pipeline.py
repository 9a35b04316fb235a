"""Host-resident batches: chunked H2D -> kernels -> D2H with the three legs overlapped.

Two entry points: `HostPipeline.run` (features + log-mel of N clips) and `TripletPipeline.run` (the training batch:
anchors and positives come from the host, the negatives are made ON THE DEVICE by the augmentation kernels from the
anchors that were just uploaded, then features + log-mel of all three -- BASELINE config 3).

The reference computes features in DataLoader workers on CPU tensors (src/data.py:231-265) and ships
stems to the GPU afterwards (src/train.py:231).  With the front end on the GPU the stems cross PCIe
first; this helper is the user-facing call for that case: pinned host stems in, pinned host features
and log-mels out, copies of chunk i+1 and i-1 overlapping the kernels of chunk i.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

from ._native import STEMS, copy2d_async


def _one_storage(tensors) -> bool:
    """True when the tensors are views of ONE allocation (a strided transfer must not cross allocations, even when separate
    pinned buffers happen to lie equally spaced)."""
    first = tensors[0].untyped_storage().data_ptr()
    return all(t.untyped_storage().data_ptr() == first for t in tensors)


class HostPipeline:
    def __init__(self, extractor, n_clips: int, T: int, chunk: int = 12, device=None, with_logmel: bool = True):
        self.fe = extractor
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        # a plan (its growable workspace) serves ONE stream at a time: the pipeline runs its kernels on a private stream,
        # so it owns a private plan instead of borrowing the extractor's
        self.plan = extractor._new_plan(self.device)
        self.n, self.T, self.chunk = n_clips, T, max(1, min(chunk, n_clips))
        self.copy2d = os.environ.get("MIXFE_PIPE_COPY2D", "1") != "0"        # strided transfers (A/B knob)
        self.with_logmel = with_logmel
        # chunk boundaries: short chunks at both ends (the first upload and the last download are the only legs
        # that nothing overlaps), full ones in between
        sizes, left, ramp = [], n_clips, [max(1, self.chunk // 4), max(1, self.chunk // 2)]
        head = [r for r in ramp if r < self.chunk]
        for r in head:
            if left > 2 * sum(head):
                sizes.append(r); left -= r
        tail = list(reversed(sizes))
        left -= sum(tail)
        while left > 0:
            m = min(self.chunk, left)
            sizes.append(m); left -= m
        sizes += tail
        self.bounds, lo = [], 0
        for m in sizes:
            self.bounds.append((lo, lo + m)); lo += m
        assert lo == n_clips
        F = self.plan.num_frames(T)
        # two device slots so that chunk i+1 uploads while chunk i computes
        self.slots = []
        for _ in range(2):
            self.slots.append({
                "flat": torch.empty((len(STEMS) * self.chunk * 2 * T,), dtype=torch.float32, device=self.device),
                "feat": torch.empty((self.chunk, self.plan.feature_dim), dtype=torch.float32, device=self.device),
                "lm": torch.empty((self.chunk, 8, self.plan.n_mels, F), dtype=torch.float32, device=self.device)
                if with_logmel else None,
                "uploaded": torch.cuda.Event(), "computed": torch.cuda.Event(), "drained": torch.cuda.Event(),
            })
        self.s_up = torch.cuda.Stream(self.device)
        self.s_run = torch.cuda.Stream(self.device)
        self.s_down = torch.cuda.Stream(self.device)
        self._n_chunks_done = 0         # chunks queued so far, over all batches: slot = parity, waits from the third on
        self.h2d_bytes = 4 * 2 * T * 4 * n_clips
        self.d2h_bytes = 4 * n_clips * (self.plan.feature_dim + (8 * self.plan.n_mels * F if with_logmel else 0))

    @staticmethod
    def pinned_like(shape) -> torch.Tensor:
        return torch.empty(shape, dtype=torch.float32, pin_memory=True)

    def run(self, host_stems: Dict[str, torch.Tensor], feat_out: torch.Tensor, lm_out: Optional[torch.Tensor]):
        """host_stems: dict of pinned (N,2,T); feat_out pinned (N,D); lm_out pinned (N,8,n_mels,F) or None.
        Returns after everything has landed in host memory."""
        self.wait(self.submit(host_stems, feat_out, lm_out))

    @staticmethod
    def wait(ticket: torch.cuda.Event) -> None:
        """Block until the batch `submit` returned the ticket for has landed in its host buffers."""
        ticket.synchronize()

    def submit(self, host_stems: Dict[str, torch.Tensor], feat_out: torch.Tensor, lm_out: Optional[torch.Tensor]):
        """`run` without the final wait: queues the whole batch and returns a ticket for `wait`.  Submitting batch
        k+1 before waiting for batch k lets its first upload run under batch k's last download (the two legs `run`
        cannot overlap); the caller then owns two sets of host output buffers, as with any prefetching loader."""
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_up, self.s_run, self.s_down):
            s.wait_stream(cur)
        for (lo, hi) in self.bounds:
            m = hi - lo
            i = self._n_chunks_done
            self._n_chunks_done += 1
            slot = self.slots[i & 1]
            per = m * 2 * self.T
            stems_d = [slot["flat"][k * per:(k + 1) * per].view(m, 2, self.T) for k in range(len(STEMS))]
            with torch.cuda.stream(self.s_up):
                if i >= 2:
                    self.s_up.wait_event(slot["computed"])       # kernels of chunk i-2 have consumed the slot
                hs = [host_stems[name] for name in STEMS]
                gap = hs[1].data_ptr() - hs[0].data_ptr()
                if self.copy2d and gap > 0 and all(h.is_contiguous() and h.is_pinned() and h.dtype == torch.float32 for h in hs) \
                        and _one_storage(hs) and all(hs[k + 1].data_ptr() - hs[k].data_ptr() == gap for k in range(len(hs) - 1)):
                    # equally spaced views of one pinned buffer: the chunk of all four stems in ONE strided transfer
                    copy2d_async(stems_d[0], per * 4, hs[0][lo], gap, per * 4, len(STEMS), self.device)
                else:
                    for d, h in zip(stems_d, hs):
                        d.copy_(h[lo:hi], non_blocking=True)
                slot["uploaded"].record(self.s_up)
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(slot["uploaded"])
                if i >= 2:
                    self.s_run.wait_event(slot["drained"])        # outputs of chunk i-2 have left the slot
                self.plan.features_logmel(stems_d, True, self.with_logmel,
                                          features_out=slot["feat"][:m],
                                          logmel_out=slot["lm"][:m] if self.with_logmel else None)
                slot["computed"].record(self.s_run)
            with torch.cuda.stream(self.s_down):
                self.s_down.wait_event(slot["computed"])
                feat_out[lo:hi].copy_(slot["feat"][:m], non_blocking=True)
                if self.with_logmel:
                    lm_out[lo:hi].copy_(slot["lm"][:m], non_blocking=True)
                slot["drained"].record(self.s_down)
        ticket = torch.cuda.Event()
        ticket.record(self.s_down)
        return ticket


class TripletPipeline:
    """The contrastive training batch from host memory (BASELINE.json config 3): G anchors and G positives cross
    PCIe, the G negatives are augment_stems(anchors) computed on the device (reference: src/data.py:231-265 calls
    AudioAugmenter in the DataLoader worker and ships three waveforms), then features + log-mel of all 3G clips.
    Outputs are ordered anchors | positives | negatives.  The extractor's and the augmenter's kernels run on this
    object's private stream, on private plans."""

    def __init__(self, extractor, augmenter, n_triplets: int, T: int, chunk: int = 4, device=None, with_logmel: bool = True):
        self.fe, self.aug = extractor, augmenter
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.plan = extractor._new_plan(self.device)
        self.G, self.T = n_triplets, T
        self.copy2d = os.environ.get("MIXFE_PIPE_COPY2D", "1") != "0"        # strided transfers (A/B knob)
        self.chunk = max(1, min(chunk, n_triplets))
        self.with_logmel = with_logmel
        sizes, left = [], n_triplets
        if n_triplets > 2 * self.chunk:                  # short first / last chunks: the legs nothing overlaps
            head = max(1, self.chunk // 2)
            sizes.append(head); left -= 2 * head
        while left > 0:
            m = min(self.chunk, left)
            sizes.append(m); left -= m
        if n_triplets > 2 * self.chunk:
            sizes.append(max(1, self.chunk // 2))
        self.bounds, lo = [], 0
        for m in sizes:
            self.bounds.append((lo, lo + m)); lo += m
        assert lo == n_triplets
        F = self.plan.num_frames(T)
        c3 = 3 * self.chunk
        self.slots = []
        for _ in range(2):
            self.slots.append({
                # one allocation for the four stems: a chunk of m triplets lays them out 3m clips apart, so that the four
                # stems of a chunk can arrive in ONE strided transfer
                "flat": torch.empty((len(STEMS) * c3 * 2 * T,), dtype=torch.float32, device=self.device),
                "feat": torch.empty((c3, self.plan.feature_dim), dtype=torch.float32, device=self.device),
                "lm": torch.empty((c3, 8, self.plan.n_mels, F), dtype=torch.float32, device=self.device) if with_logmel else None,
                "uploaded": torch.cuda.Event(), "computed": torch.cuda.Event(), "drained": torch.cuda.Event(),
            })
        self.s_up = torch.cuda.Stream(self.device)
        self.s_run = torch.cuda.Stream(self.device)
        self.s_down = torch.cuda.Stream(self.device)
        self._n_chunks_done = 0
        self.h2d_bytes = 4 * 2 * T * 4 * 2 * n_triplets
        self.d2h_bytes = 4 * 3 * n_triplets * (self.plan.feature_dim + (8 * self.plan.n_mels * F if with_logmel else 0))

    def run(self, host_stems: Dict[str, torch.Tensor], decisions, feat_out: torch.Tensor, lm_out: Optional[torch.Tensor]):
        """host_stems: dict of pinned (2G,2,T), anchors then positives; decisions: G augmentation decisions (a list of
        ClipAugmentation or the BatchAugmentation of AudioAugmenter.draw_batch); feat_out pinned (3G,D); lm_out pinned
        (3G,8,n_mels,F) or None.  Returns after everything has landed in host memory.  feat_out / lm_out may also be CUDA
        tensors on the pipeline's device: the results then stay on the GPU (the encoder's input) and only the stems cross PCIe."""
        self.wait(self.submit(host_stems, decisions, feat_out, lm_out))

    wait = staticmethod(HostPipeline.wait)

    def submit(self, host_stems: Dict[str, torch.Tensor], decisions, feat_out: torch.Tensor, lm_out: Optional[torch.Tensor]):
        """`run` without the final wait (see HostPipeline.submit): returns a ticket for `wait`."""
        G = self.G
        if len(decisions) != G:
            raise ValueError(f"{len(decisions)} decisions for {G} anchors")
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_up, self.s_run, self.s_down):
            s.wait_stream(cur)
        for (lo, hi) in self.bounds:
            m = hi - lo
            i = self._n_chunks_done
            self._n_chunks_done += 1
            slot = self.slots[i & 1]
            per = 3 * m * 2 * self.T
            stems_d = [slot["flat"][k * per:(k + 1) * per].view(3 * m, 2, self.T) for k in range(len(STEMS))]
            with torch.cuda.stream(self.s_up):
                if i >= 2:
                    self.s_up.wait_event(slot["computed"])
                hs = [host_stems[name] for name in STEMS]
                row = 2 * self.T * 4
                strided = self.copy2d and all(h.is_contiguous() and h.is_pinned() and h.dtype == torch.float32 for h in hs)
                gap = hs[1].data_ptr() - hs[0].data_ptr()
                if strided and gap > 0 and gap % (G * row) == 0 and _one_storage(hs) and \
                        all(hs[k + 1].data_ptr() - hs[k].data_ptr() == gap for k in range(len(hs) - 1)):
                    # the caller's four stem tensors are equally spaced views of one pinned allocation: anchors and positives
                    # of all four stems in ONE transfer (blocks = stems, rows = anchors | positives)
                    copy2d_async(stems_d[0], m * row, hs[0][lo], G * row, m * row, 2, self.device,
                                 blocks=len(STEMS), dst_block_rows=3, src_block_rows=gap // (G * row))
                elif strided:
                    # anchors [lo, hi) and positives [G + lo, G + hi): two runs one stride apart -> one transfer per stem
                    for d, h in zip(stems_d, hs):
                        copy2d_async(d, m * row, h[lo], G * row, m * row, 2, self.device)
                else:
                    for d, h in zip(stems_d, hs):
                        d[:m].copy_(h[lo:hi], non_blocking=True)
                        d[m:2 * m].copy_(h[G + lo:G + hi], non_blocking=True)
                slot["uploaded"].record(self.s_up)
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(slot["uploaded"])
                if i >= 2:
                    self.s_run.wait_event(slot["drained"])
                dec = decisions.slice(lo, hi) if hasattr(decisions, "slice") else decisions[lo:hi]
                self.aug.apply({n: d[:m] for n, d in zip(STEMS, stems_d)}, dec,
                               out={n: d[2 * m:3 * m] for n, d in zip(STEMS, stems_d)})
                self.plan.features_logmel(stems_d, True, self.with_logmel,
                                          features_out=slot["feat"][:3 * m],
                                          logmel_out=slot["lm"][:3 * m] if self.with_logmel else None)
                slot["computed"].record(self.s_run)
            with torch.cuda.stream(self.s_down):
                self.s_down.wait_event(slot["computed"])
                if self.copy2d and feat_out.is_contiguous() and feat_out.is_pinned() and \
                        (lm_out is None or (lm_out.is_contiguous() and lm_out.is_pinned())):
                    # the three parts (anchors | positives | negatives) of the chunk land G rows apart: one transfer each
                    rf = feat_out.shape[1] * 4
                    copy2d_async(feat_out[lo], G * rf, slot["feat"], m * rf, m * rf, 3, self.device)
                    if self.with_logmel:
                        rl = lm_out[0].numel() * 4
                        copy2d_async(lm_out[lo], G * rl, slot["lm"], m * rl, m * rl, 3, self.device)
                else:
                    for part in range(3):
                        feat_out[part * G + lo:part * G + hi].copy_(slot["feat"][part * m:(part + 1) * m], non_blocking=True)
                        if self.with_logmel:
                            lm_out[part * G + lo:part * G + hi].copy_(slot["lm"][part * m:(part + 1) * m], non_blocking=True)
                slot["drained"].record(self.s_down)
        ticket = torch.cuda.Event()
        ticket.record(self.s_down)
        return ticket
