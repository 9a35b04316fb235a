"""GPU: mixfe_weighted_masking_noalloc (include/mixfe.h) -- the allocation-free variant of mixfe_weighted_masking whose
scratch is a static __device__ array inside libmixfe -- called THROUGH THE C ABI (ctypes, raw pointers), against the
existing entry point on the same inputs, on two calls in a row (the second on a different, smaller batch: whatever the
first call left in the shared array must not matter).  Tolerance: that of tests/test_gpu_extensions.py (2e-5 absolute on
values in [0, 1]); the two entry points run the same kernels, so the results are in fact equal bit for bit.

Collected last (test_zz_*)."""
import ctypes

import pytest
import torch

import mixing_style_transfer_b200 as pkg
from mixing_style_transfer_b200 import _native
from oracle import mixfe_oracle as O

pytestmark = pytest.mark.gpu

TOL = 2e-5
MAX_BLOCKS = 8192            # MIXFE_NOALLOC_MASKING_MAX_BLOCKS of include/mixfe.h


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _logmel(fe, dev, seeds, T):
    clips = [O.musical_clip(s, T) for s in seeds]
    batch = {n: torch.stack([c[n] for c in clips]).to(dev) for n in O.STEMS}
    _, lm = fe.features_and_logmel(batch)
    return lm.contiguous()


def _both(plan, lm, T, bw, bw_sum, beta, tau, dev):
    """(existing entry point, no-alloc entry point) on the same log-mel, both through ctypes."""
    L = _native.lib()
    B = lm.shape[0]
    out_ref = torch.full((B, 4), -1.0, device=dev)
    out_new = torch.full((B, 4), -2.0, device=dev)
    args = (ctypes.c_void_p(lm.data_ptr()), lm.stride(0), B, T, ctypes.c_void_p(bw.data_ptr()), bw_sum, beta, tau)
    with torch.cuda.device(dev):
        rc = L.mixfe_weighted_masking(plan._h, *args, ctypes.c_void_p(out_ref.data_ptr()), _stream(dev))
        assert rc == 0, L.mixfe_last_error()
        rc = L.mixfe_weighted_masking_noalloc(plan._h, *args, ctypes.c_void_p(out_new.data_ptr()), _stream(dev))
        assert rc == 0, L.mixfe_last_error()
    torch.cuda.synchronize(dev)
    return out_ref, out_new


def test_noalloc_masking_matches_the_existing_entry_point_twice(cuda_device):
    dev = cuda_device
    fe = pkg.MixingFeatureExtractor()
    plan = fe._plan(dev)
    bw = fe.band_weights().to(dev, torch.float32).contiguous()
    bw_sum = float(fe.band_weights().double().sum())
    T1, T2 = 40000, 23000
    lm1 = _logmel(fe, dev, (60, 61, 62, 63, 64), T1)
    lm2 = _logmel(fe, dev, (70, 71), T2)
    mem0 = torch.cuda.memory_allocated(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    ref1, new1 = _both(plan, lm1, T1, bw, bw_sum, 0.0, 1.0, dev)
    ref2, new2 = _both(plan, lm2, T2, bw, bw_sum, 0.5, 0.25, dev)      # second call in a row: smaller batch, other constants
    for ref, new in ((ref1, new1), (ref2, new2)):
        assert bool(((ref >= 0) & (ref <= 1)).all()) and float(ref.abs().max()) > 0.0
        err = float((ref - new).abs().max())
        print(f"no-alloc vs existing entry point: max abs diff {err:.3g}")
        assert err <= TOL
    # and a third, the first batch again: identical to its first result
    _, again = _both(plan, lm1, T1, bw, bw_sum, 0.0, 1.0, dev)
    assert torch.equal(again, new1)
    # the new entry point itself allocated nothing: repeat it alone and watch the driver's free memory
    L = _native.lib()
    out = torch.empty((lm1.shape[0], 4), device=dev)
    torch.cuda.synchronize(dev)
    free1, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        with torch.cuda.device(dev):
            assert L.mixfe_weighted_masking_noalloc(plan._h, ctypes.c_void_p(lm1.data_ptr()), lm1.stride(0), lm1.shape[0], T1,
                                                    ctypes.c_void_p(bw.data_ptr()), bw_sum, 0.0, 1.0,
                                                    ctypes.c_void_p(out.data_ptr()), _stream(dev)) == 0
    torch.cuda.synchronize(dev)
    free2, _ = torch.cuda.mem_get_info(dev)
    assert free2 == free1
    assert torch.equal(out, new1)
    del mem0, free0


def test_noalloc_masking_rejects_sizes_beyond_the_static_scratch(cuda_device):
    dev = cuda_device
    fe = pkg.MixingFeatureExtractor()
    plan = fe._plan(dev)
    L = _native.lib()
    T = 441000                                   # 1723 frames = 14 blocks of 128 per clip
    blocks = -(-(1 + T // 256) // 128)
    B = MAX_BLOCKS // blocks + 1                 # one clip too many; the size check precedes any memory access
    dummy = torch.zeros(16, device=dev)
    out = torch.zeros((4,), device=dev)
    with torch.cuda.device(dev):
        rc = L.mixfe_weighted_masking_noalloc(plan._h, ctypes.c_void_p(dummy.data_ptr()), 8 * 128 * (1 + T // 256), B, T,
                                              ctypes.c_void_p(dummy.data_ptr()), 1.0, 0.0, 1.0,
                                              ctypes.c_void_p(out.data_ptr()), _stream(dev))
    assert rc == _native.E_UNSUPPORTED
    msg = L.mixfe_last_error().decode()
    assert "MIXFE_NOALLOC_MASKING_MAX_BLOCKS" in msg and str(MAX_BLOCKS) in msg
    # an augmentation-only plan has no feature tables
    aug_plan = pkg.AudioAugmenter()._plan(dev)
    with torch.cuda.device(dev):
        rc = L.mixfe_weighted_masking_noalloc(aug_plan._h, ctypes.c_void_p(dummy.data_ptr()), 0, 1, 4000,
                                              ctypes.c_void_p(dummy.data_ptr()), 1.0, 0.0, 1.0,
                                              ctypes.c_void_p(out.data_ptr()), _stream(dev))
    assert rc == _native.E_INVAL
