"""ctypes binding of libmixfe.so (include/mixfe.h) and its in-tree nvcc build.

There is NO CPU fallback: if the shared library is missing or cannot be loaded, every entry point
raises.  Tensors are only used here as owners of device memory; the arithmetic is in csrc/.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading
from typing import List, Optional, Sequence

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_PKG)
CSRC = os.path.join(_PKG, "csrc")
# MIXFE_LIB: developer knob for A/B runs of kernel variants (scripts/): load another build of the same ABI
LIB_PATH = os.environ.get("MIXFE_LIB") or os.path.join(_PKG, "libmixfe.so")
INCLUDE = os.path.join(_REPO, "include", "mixfe.h")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]

STEMS = ("vocals", "bass", "drums", "other")   # reference: src/model.py:54, src/mixing_utils.py:287

E_INVAL, E_UNSUPPORTED, E_CUDA, E_SHORT = -1, -2, -3, -4
ABI_VERSION = 2                                   # MIXFE_ABI_VERSION of include/mixfe.h


def _sources():
    out = [INCLUDE]
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cu", ".cuh", ".h")):
            out.append(os.path.join(CSRC, f))
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/mixfe.cu for sm_100a into libmixfe.so next to this file (needs nvcc, no GPU)."""
    newest = max(os.path.getmtime(p) for p in _sources())
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= newest:
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB_PATH, os.path.join(CSRC, "mixfe.cu")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB_PATH


class AugParams(ctypes.Structure):
    """mixfe_aug_params (include/mixfe.h)."""
    _fields_ = [
        ("gain_lin", ctypes.c_float * 4),
        ("tilt_on", ctypes.c_int * 4),
        ("tilt_sos", (ctypes.c_double * 6) * 4),
        ("comp_on", ctypes.c_int * 4),
        ("comp_threshold_db", ctypes.c_float * 4),
        ("comp_ratio", ctypes.c_float * 4),
        ("bw_on", ctypes.c_int * 4),
        ("bw_sos", (ctypes.c_double * 12) * 4),
        ("reverb_on", ctypes.c_int),
        ("_pad", ctypes.c_int),
    ]


_lib: Optional[ctypes.CDLL] = None
_lock = threading.Lock()

# every symbol include/mixfe.h declares (tests check the library exports all of them)
EXPORTS = ("mixfe_abi_version", "mixfe_last_error", "mixfe_plan_create", "mixfe_plan_destroy",
           "mixfe_plan_set_filterbank", "mixfe_feature_dim", "mixfe_num_frames", "mixfe_workspace_bytes",
           "mixfe_features_logmel", "mixfe_logmel_backward", "mixfe_augment", "mixfe_debug_mel_power", "mixfe_launch_count",
           "mixfe_plan_enable_timing", "mixfe_plan_kernel_time", "mixfe_debug_feature_kernel_config",
           "mixfe_plan_reserve_sms", "mixfe_copy_async", "mixfe_gated_loudness", "mixfe_short_term_loudness_iqr",
           "mixfe_stft_loss_forward", "mixfe_stft_loss_backward", "mixfe_plan_create_augment", "mixfe_effects",
           "mixfe_reverb", "mixfe_weighted_masking", "mixfe_envelope_compress", "mixfe_push_peers",
           "mixfe_weighted_masking_noalloc", "mixfe_copy_strided_async")


def lib() -> ctypes.CDLL:
    """Load libmixfe.so (once).  Raises RuntimeError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        L.mixfe_abi_version.restype = ctypes.c_int
        if L.mixfe_abi_version() != ABI_VERSION:
            raise RuntimeError(f"{LIB_PATH} has ABI version {L.mixfe_abi_version()}, this package binds version "
                               f"{ABI_VERSION}: rebuild it (`python -c 'import __graft_entry__ as g; g.build()'`)")
        vp, ll, ci = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int
        L.mixfe_abi_version.restype = ci
        L.mixfe_last_error.restype = ctypes.c_char_p
        L.mixfe_plan_create.argtypes = [ctypes.POINTER(vp), ci, ci, ci, ci, ci, ci]
        L.mixfe_plan_destroy.argtypes = [vp]
        L.mixfe_plan_set_filterbank.argtypes = [vp, vp]
        L.mixfe_feature_dim.argtypes = [vp]
        L.mixfe_num_frames.argtypes = [vp, ci]
        L.mixfe_workspace_bytes.argtypes = [vp]
        L.mixfe_workspace_bytes.restype = ctypes.c_size_t
        L.mixfe_features_logmel.argtypes = [vp, ctypes.POINTER(vp), ll, ll, ci, ci, vp, vp, ll, vp]
        L.mixfe_debug_mel_power.argtypes = [vp, ctypes.POINTER(vp), ll, ll, ci, ci, vp, ll, vp]
        L.mixfe_logmel_backward.argtypes = [vp, ctypes.POINTER(vp), ll, ll, ci, ci, vp, ll, vp, ll, ctypes.POINTER(vp), ll, ll, vp]
        L.mixfe_augment.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), ll, ll, ci, ci,
                                    ctypes.POINTER(AugParams), vp, ci, vp]
        L.mixfe_plan_create_augment.argtypes = [ctypes.POINTER(vp), ci, ci]
        L.mixfe_effects.argtypes = [vp, vp, vp, ll, ll, ci, ci, ci, ctypes.POINTER(AugParams), vp]
        L.mixfe_reverb.argtypes = [vp, vp, vp, ll, ll, ci, ci, ci, vp, ll, ci, vp]
        L.mixfe_weighted_masking.argtypes = [vp, vp, ll, ci, ci, vp, ctypes.c_double, ctypes.c_float, ctypes.c_float, vp, vp]
        L.mixfe_weighted_masking_noalloc.argtypes = L.mixfe_weighted_masking.argtypes
        L.mixfe_envelope_compress.argtypes = [vp, vp, vp, ll, ll, ci, ci, ci, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                                              ctypes.c_float, vp]
        L.mixfe_launch_count.restype = ll
        L.mixfe_debug_feature_kernel_config.argtypes = [vp, ctypes.POINTER(ci), ctypes.POINTER(ci), ctypes.POINTER(ci)]
        L.mixfe_plan_reserve_sms.argtypes = [vp, ci]
        L.mixfe_copy_async.argtypes = [vp, vp, ctypes.c_size_t, vp]
        sz = ctypes.c_size_t
        L.mixfe_copy_strided_async.argtypes = [vp, sz, sz, vp, sz, sz, sz, sz, sz, vp]
        L.mixfe_push_peers.argtypes = [ctypes.POINTER(vp), ci, vp, ctypes.c_size_t, ci, vp]
        L.mixfe_gated_loudness.argtypes = [vp, ctypes.POINTER(vp), ll, ll, ci, ci, vp, vp]
        L.mixfe_short_term_loudness_iqr.argtypes = [vp, ctypes.POINTER(vp), ll, ll, ci, ci, vp, vp]
        L.mixfe_stft_loss_forward.argtypes = [vp, vp, vp, ci, ci, ci, ll, ll, ll, ll, vp, vp]
        L.mixfe_stft_loss_backward.argtypes = [vp, vp, vp, ci, ci, ci, ll, ll, ll, ll, vp, vp, vp, ll, ll, vp]
        L.mixfe_plan_enable_timing.argtypes = [vp, ci]
        L.mixfe_plan_kernel_time.argtypes = [vp, ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ll)]
        _lib = L
    return _lib


# Tracing hook: MIXFE_NVTX=1 brackets every entry point of a Plan with an NVTX range
# ("mixfe.features_logmel", "mixfe.augment", ...), so an nsys / ncu --nvtx timeline shows the front end's calls between the
# caller's own ranges.  Off by default: the range push/pop is host work on the launch path.
NVTX = os.environ.get("MIXFE_NVTX", "0") not in ("", "0")


def _traced(fn):
    if not NVTX:
        return fn
    import functools
    name = "mixfe." + fn.__name__

    @functools.wraps(fn)
    def wrapper(*a, **k):
        torch.cuda.nvtx.range_push(name)
        try:
            return fn(*a, **k)
        finally:
            torch.cuda.nvtx.range_pop()
    return wrapper


def launch_count() -> int:
    return int(lib().mixfe_launch_count())


def _check(rc: int) -> None:
    if rc == 0:
        return
    msg = lib().mixfe_last_error().decode("utf-8", "replace")
    if rc == E_SHORT:
        raise RuntimeError(msg)                 # what torch.stft's reflect pad raises in the reference
    if rc == E_INVAL:
        raise ValueError(msg)
    if rc == E_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(f"libmixfe error {rc}: {msg}")


def _ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _stream(device: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def unify_stems(stems: Sequence[torch.Tensor]):
    """4 x (B,2,T) CUDA fp32 tensors with unit time stride and a COMMON (batch, channel) stride."""
    st = [s if (s.dtype == torch.float32) else s.float() for s in stems]
    shape = st[0].shape
    for s in st:
        if s.shape != shape:
            raise ValueError(f"stems have different shapes: {tuple(s.shape)} vs {tuple(shape)}")
    strides = {(s.stride(0), s.stride(1), s.stride(2)) for s in st}
    ok = len(strides) == 1 and st[0].stride(2) == 1 and st[0].stride(1) >= shape[2] and \
        (shape[0] == 1 or st[0].stride(0) >= shape[2])
    if not ok:
        st = [s.contiguous() for s in st]
    return st, st[0].stride(0), st[0].stride(1)


def _check_out(t: torch.Tensor, shape, device: torch.device, name: str) -> None:
    """A caller-provided output buffer must be fp32, on the plan's device and exactly `shape`; strides are checked by
    the caller where the kernel assumes a layout.  Raises ValueError instead of letting a kernel write out of bounds."""
    if t.dtype != torch.float32:
        raise ValueError(f"{name}: dtype {t.dtype}, expected torch.float32")
    if t.device != device:
        raise ValueError(f"{name}: on {t.device}, the plan lives on {device}")
    if tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name}: shape {tuple(t.shape)}, expected {tuple(shape)}")


def _span(t: torch.Tensor):
    """[first byte, last byte + 1) of the memory a strided tensor touches."""
    if t.numel() == 0:
        return (t.data_ptr(), t.data_ptr())
    last = sum((n - 1) * st for n, st in zip(t.shape, t.stride()))
    return (t.data_ptr(), t.data_ptr() + (last + 1) * t.element_size())


def copy_async(dst_ptr: int, src: torch.Tensor, nbytes: int) -> None:
    """Copy-engine D2D copy of `nbytes` from the (contiguous) tensor `src` to the raw device address `dst_ptr` on the
    current stream of src's device; dst_ptr may be an NVSwitch multicast address."""
    _check(lib().mixfe_copy_async(ctypes.c_void_p(dst_ptr), ctypes.c_void_p(src.data_ptr()), nbytes, _stream(src.device)))


def copy2d_async(dst: torch.Tensor, dst_pitch: int, src: torch.Tensor, src_pitch: int, width_bytes: int, rows: int,
                 device, blocks: int = 1, dst_block_rows: int = 0, src_block_rows: int = 0) -> None:
    """Copy-engine strided copy on the current stream of `device`: `blocks` x `rows` runs of width_bytes starting at dst / src
    (tensor views whose data_ptr is the first byte), pitches in bytes, block strides in rows; host tensors must be pinned for
    the copy to be asynchronous."""
    _check(lib().mixfe_copy_strided_async(ctypes.c_void_p(dst.data_ptr()), dst_pitch, dst_block_rows,
                                          ctypes.c_void_p(src.data_ptr()), src_pitch, src_block_rows,
                                          width_bytes, rows, blocks, _stream(device)))


def push_peers(dst_ptrs: Sequence[int], src: torch.Tensor, nbytes: int, n_ctas: int) -> None:
    """SM-driven copy of `nbytes` from the (contiguous) tensor `src` to every raw device address in dst_ptrs (NVLink peer
    mappings) on the current stream of src's device (kernel K11)."""
    arr = (ctypes.c_void_p * len(dst_ptrs))(*dst_ptrs)
    _check(lib().mixfe_push_peers(arr, len(dst_ptrs), ctypes.c_void_p(src.data_ptr()), nbytes, n_ctas, _stream(src.device)))


class Plan:
    """Owner of a mixfe_plan (constant tables + workspace) bound to one CUDA device."""

    def __init__(self, device: torch.device, sample_rate: int, n_fft: int = 0, hop: int = 0, n_mels: int = 0,
                 n_spectral_bins: int = 0, fb: Optional[torch.Tensor] = None):
        """n_fft = 0: an augmentation-only plan (mixfe_plan_create_augment), no mel / STFT tables."""
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("libmixfe plans live on CUDA devices; there is no CPU path")
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        torch.cuda.init()
        with torch.cuda.device(idx):
            torch.cuda.current_stream()          # make sure the primary context exists
            h = ctypes.c_void_p()
            if n_fft:
                _check(lib().mixfe_plan_create(ctypes.byref(h), idx, sample_rate, n_fft, hop, n_mels, n_spectral_bins))
            else:
                _check(lib().mixfe_plan_create_augment(ctypes.byref(h), idx, sample_rate))
        self._h = h
        self.n_mels = n_mels
        self.hop = hop
        self.n_fft = n_fft
        self.feature_dim = int(lib().mixfe_feature_dim(self._h)) if n_fft else 0
        if fb is not None:
            self.set_filterbank(fb)

    def set_filterbank(self, fb: torch.Tensor) -> None:
        fb = fb.detach().to("cpu", torch.float32).contiguous()
        if tuple(fb.shape) != (self.n_fft // 2 + 1, self.n_mels):
            raise ValueError(f"filterbank shape {tuple(fb.shape)}")
        _check(lib().mixfe_plan_set_filterbank(self._h, ctypes.c_void_p(fb.data_ptr())))

    def num_frames(self, T: int) -> int:
        if not self.n_fft:
            raise ValueError("augmentation-only plan (AudioAugmenter's): it has no STFT / mel tables")
        return 1 + T // self.hop

    def workspace_bytes(self) -> int:
        return int(lib().mixfe_workspace_bytes(self._h))

    def k1_config(self) -> dict:
        """Launch shape of the fused feature kernel: resident CTAs per SM, dynamic smem, mel-walk steps."""
        a, b, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _check(lib().mixfe_debug_feature_kernel_config(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return {"ctas_per_sm": a.value, "smem_bytes": b.value, "mel_steps": c.value}

    @_traced
    def features_logmel(self, stems: Sequence[torch.Tensor], want_features: bool = True,
                        want_logmel: bool = True, features_out: Optional[torch.Tensor] = None,
                        logmel_out: Optional[torch.Tensor] = None):
        """stems: 4 CUDA tensors (B,2,T).  Returns (features (B,D) | None, logmel (B,8,n_mels,F) | None)."""
        st, sb, sc = unify_stems(stems)
        B, _, T = st[0].shape
        F = self.num_frames(T)
        feats = lm = None
        if want_features:
            feats = features_out if features_out is not None else \
                torch.empty((B, self.feature_dim), dtype=torch.float32, device=self.device)
            if features_out is not None:
                _check_out(feats, (B, self.feature_dim), self.device, "features_out")
                if B and (feats.stride(1) != 1 or (B > 1 and feats.stride(0) != self.feature_dim)):
                    raise ValueError("features_out must be a dense (B, D) buffer")
        if want_logmel:
            lm = logmel_out if logmel_out is not None else \
                torch.empty((B, 8, self.n_mels, F), dtype=torch.float32, device=self.device)
            if logmel_out is not None:
                _check_out(lm, (B, 8, self.n_mels, F), self.device, "logmel_out")
            if lm.stride(3) != 1 or lm.stride(2) != F or lm.stride(1) != self.n_mels * F or \
                    (B > 1 and lm.stride(0) < 8 * self.n_mels * F):
                raise ValueError("logmel_out must be contiguous in its last three dims (any batch stride >= one clip)")
        ptrs = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in st])
        with torch.cuda.device(self.device):
            _check(lib().mixfe_features_logmel(self._h, ptrs, sb, sc, B, T, _ptr(feats), _ptr(lm),
                                               0 if lm is None else lm.stride(0), _stream(self.device)))
        return feats, lm

    @_traced
    def logmel_backward(self, stems: Sequence[torch.Tensor], grad_logmel: torch.Tensor,
                        logmel: Optional[torch.Tensor] = None) -> List[torch.Tensor]:
        """d loss / d stems (4 new (B,2,T) tensors) from d loss / d logmel (B,8,n_mels,F); the adjoint of the log-mel half
        of features_logmel.  `logmel`: the forward's output for the same stems -- with it the (1024, 256) configuration
        runs K6F (one write per gradient sample, no mel recomputation), else the generic K6."""
        st, sb, sc = unify_stems(stems)
        B, _, T = st[0].shape
        F = self.num_frames(T)

        def dense(x, name):
            if x.dtype != torch.float32 or x.device != self.device:
                x = x.to(device=self.device, dtype=torch.float32)
            if tuple(x.shape) != (B, 8, self.n_mels, F):
                raise ValueError(f"{name} shape {tuple(x.shape)}, expected {(B, 8, self.n_mels, F)}")
            if x.stride(3) != 1 or x.stride(2) != F or x.stride(1) != self.n_mels * F:
                x = x.contiguous()
            return x
        g = dense(grad_logmel, "grad_logmel")
        lm = dense(logmel, "logmel") if logmel is not None else None
        outs = [torch.empty((B, 2, T), dtype=torch.float32, device=self.device) for _ in range(4)]
        ptrs = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in st])
        optrs = (ctypes.c_void_p * 4)(*[o.data_ptr() for o in outs])
        with torch.cuda.device(self.device):
            _check(lib().mixfe_logmel_backward(self._h, ptrs, sb, sc, B, T, _ptr(lm), 0 if lm is None else lm.stride(0),
                                               _ptr(g), g.stride(0), optrs, outs[0].stride(0), outs[0].stride(1),
                                               _stream(self.device)))
        return outs

    @staticmethod
    def _bct(t: torch.Tensor) -> torch.Tensor:
        if t.dtype != torch.float32:
            t = t.float()
        return t if t.stride(2) == 1 else t.contiguous()

    @_traced
    def stft_loss_forward(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """One resolution (this plan's n_fft, hop) of the STFT loss on (B,C,T) CUDA tensors -> state (8 doubles on the
        device): [3] spectral convergence, [4] log-magnitude L1 (kernel K8)."""
        x, y = self._bct(x), self._bct(y)
        B, C, T = x.shape
        state = torch.empty(8, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().mixfe_stft_loss_forward(self._h, _ptr(x), _ptr(y), B, C, T, x.stride(0), x.stride(1), y.stride(0),
                                                 y.stride(1), _ptr(state), _stream(self.device)))
        return state

    @_traced
    def stft_loss_backward(self, x: torch.Tensor, y: torch.Tensor, state: torch.Tensor, grad_scale: torch.Tensor,
                           grad_x: torch.Tensor) -> None:
        """Adds grad_scale * d(sc + lm)/dx of this resolution into grad_x ((B,C,T) fp32, contiguous in T)."""
        x, y = self._bct(x), self._bct(y)
        B, C, T = x.shape
        _check_out(grad_x, (B, C, T), self.device, "grad_x")
        if grad_x.stride(2) != 1 or grad_x.stride(1) < T or (B > 1 and grad_x.stride(0) < T):
            raise ValueError("grad_x must have unit stride in T and non-overlapping rows")
        with torch.cuda.device(self.device):
            _check(lib().mixfe_stft_loss_backward(self._h, _ptr(x), _ptr(y), B, C, T, x.stride(0), x.stride(1), y.stride(0),
                                                  y.stride(1), _ptr(state), _ptr(grad_scale), _ptr(grad_x), grad_x.stride(0),
                                                  grad_x.stride(1), _stream(self.device)))

    @_traced
    def gated_loudness(self, stems: Sequence[torch.Tensor]) -> torch.Tensor:
        """(B,5) BS.1770-4 integrated loudness [vocals, bass, drums, other, mixture] in LUFS (kernel K7; extension)."""
        st, sb, sc = unify_stems(stems)
        B, _, T = st[0].shape
        out = torch.empty((B, 5), dtype=torch.float32, device=self.device)
        ptrs = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in st])
        with torch.cuda.device(self.device):
            _check(lib().mixfe_gated_loudness(self._h, ptrs, sb, sc, B, T, _ptr(out), _stream(self.device)))
        return out

    @_traced
    def short_term_loudness_iqr(self, stems: Sequence[torch.Tensor]) -> torch.Tensor:
        """(B,5) inter-quartile range of the short-term (3 s / 100 ms) K-weighted loudness in LU, same order (kernel K7;
        extension: the proposal's DR_IQR)."""
        st, sb, sc = unify_stems(stems)
        B, _, T = st[0].shape
        out = torch.empty((B, 5), dtype=torch.float32, device=self.device)
        ptrs = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in st])
        with torch.cuda.device(self.device):
            _check(lib().mixfe_short_term_loudness_iqr(self._h, ptrs, sb, sc, B, T, _ptr(out), _stream(self.device)))
        return out

    def debug_mel_power(self, stems: Sequence[torch.Tensor]) -> torch.Tensor:
        st, sb, sc = unify_stems(stems)
        B, _, T = st[0].shape
        out = torch.empty((B, 8, self.n_mels, self.num_frames(T)), dtype=torch.float32, device=self.device)
        ptrs = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in st])
        with torch.cuda.device(self.device):
            _check(lib().mixfe_debug_mel_power(self._h, ptrs, sb, sc, B, T, _ptr(out), out.stride(0),
                                               _stream(self.device)))
        return out

    @_traced
    def augment(self, stems: Sequence[torch.Tensor], params, impulse: Optional[torch.Tensor],
                outs: Optional[Sequence[torch.Tensor]] = None):
        """stems: 4 CUDA tensors (B,2,T); params: ctypes array of B AugParams; impulse (B,L) or None.
        outs: optional 4 preallocated CUDA tensors with the inputs' shape AND strides (the C ABI addresses
        input and output with one pair of strides), not aliasing the inputs."""
        st, sb, sc = unify_stems(stems)
        B, _, T = st[0].shape
        if outs is None:
            outs = [torch.empty_strided(s.shape, s.stride(), dtype=torch.float32, device=self.device) for s in st]
        else:
            outs = list(outs)
            for o, s in zip(outs, st):
                if (o.shape != s.shape or o.stride() != s.stride() or o.dtype != torch.float32
                        or o.device != s.device):
                    raise ValueError("augment: `outs` must match the inputs' shape, strides, dtype and device")
        spans_in = [_span(t) for t in st]
        for o in outs:
            lo, hi = _span(o)
            if any(lo < b and a < hi for a, b in spans_in):
                raise ValueError("augment: an output buffer overlaps an input (the kernels read ahead of what they write)")
        pin = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in st])
        pout = (ctypes.c_void_p * 4)(*[s.data_ptr() for s in outs])
        L = 0
        if impulse is not None:
            impulse = impulse.to(self.device, torch.float32).contiguous()
            L = impulse.shape[-1]
        with torch.cuda.device(self.device):
            _check(lib().mixfe_augment(self._h, pin, pout, sb, sc, B, T, params, _ptr(impulse), L,
                                       _stream(self.device)))
        return outs

    @_traced
    def effects(self, x: torch.Tensor, params, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x: CUDA (B,C,T); params: ctypes array of B AugParams whose stem index 0 holds the decisions (gain, tilt,
        compressor, band-limit) applied to every channel of clip b.  Returns a new tensor (or `out`)."""
        if x.dtype != torch.float32:
            x = x.float()
        if x.stride(2) != 1:
            x = x.contiguous()
        B, C, T = x.shape
        if out is None:
            out = torch.empty_strided(x.shape, x.stride(), dtype=torch.float32, device=self.device)
        else:
            _check_out(out, x.shape, self.device, "out")
            if out.stride() != x.stride():
                raise ValueError("effects: `out` must have the input's strides")
            (a, b_), (c, d) = _span(x), _span(out)
            if c < b_ and a < d:
                raise ValueError("effects: the output buffer overlaps the input")
        with torch.cuda.device(self.device):
            _check(lib().mixfe_effects(self._h, _ptr(x), _ptr(out), x.stride(0), x.stride(1), B, C, T, params,
                                       _stream(self.device)))
        return out

    @_traced
    def reverb(self, x: torch.Tensor, impulse: torch.Tensor) -> torch.Tensor:
        """x: CUDA (B,C,T), C in {1, 2}; impulse (L,) shared by all clips or (B,L).  0.7 x + 0.3 corr(x, h) (kernel K4)."""
        if x.dtype != torch.float32:
            x = x.float()
        if x.stride(2) != 1:
            x = x.contiguous()
        B, C, T = x.shape
        impulse = impulse.to(self.device, torch.float32).contiguous()
        if impulse.dim() == 1:
            stride, L = 0, impulse.shape[0]
        elif impulse.dim() == 2 and impulse.shape[0] == B:
            stride, L = impulse.stride(0), impulse.shape[1]
        else:
            raise ValueError(f"impulse shape {tuple(impulse.shape)}: (L,) or ({B}, L)")
        out = torch.empty_strided(x.shape, x.stride(), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().mixfe_reverb(self._h, _ptr(x), _ptr(out), x.stride(0), x.stride(1), B, C, T, _ptr(impulse),
                                      stride, L, _stream(self.device)))
        return out

    @_traced
    def weighted_masking(self, logmel: torch.Tensor, T: int, band_weights: torch.Tensor, beta: float, tau: float) -> torch.Tensor:
        """(B,4) activity- and band-weighted masking from the log-mel tensor of the same clips (kernel K9; extension)."""
        B = logmel.shape[0]
        F = self.num_frames(T)
        _check_out(logmel, (B, 8, self.n_mels, F), self.device, "logmel")
        if logmel.stride(3) != 1 or logmel.stride(2) != F or logmel.stride(1) != self.n_mels * F:
            logmel = logmel.contiguous()
        bw = band_weights.to(self.device, torch.float32).contiguous()
        if tuple(bw.shape) != (self.n_mels,):
            raise ValueError(f"band_weights shape {tuple(bw.shape)}, expected ({self.n_mels},)")
        out = torch.empty((B, 4), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().mixfe_weighted_masking(self._h, _ptr(logmel), logmel.stride(0) if B else 0, B, T, _ptr(bw),
                                                float(band_weights.double().sum()), float(beta), float(tau), _ptr(out),
                                                _stream(self.device)))
        return out

    @_traced
    def envelope_compress(self, x: torch.Tensor, threshold_db: float, ratio: float, attack_coeff: float,
                          release_coeff: float) -> torch.Tensor:
        """x: CUDA (B,C,T) -> envelope-follower compressor output (kernel K10; extension)."""
        if x.dtype != torch.float32:
            x = x.float()
        if x.stride(2) != 1:
            x = x.contiguous()
        B, C, T = x.shape
        out = torch.empty_strided(x.shape, x.stride(), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().mixfe_envelope_compress(self._h, _ptr(x), _ptr(out), x.stride(0), x.stride(1), B, C, T,
                                                 float(threshold_db), float(ratio), float(attack_coeff), float(release_coeff),
                                                 _stream(self.device)))
        return out

    KERNELS = {"features": 0, "finalize": 1, "augment": 2, "reverb": 3, "logmel_bwd": 4, "loudness": 5, "stft_loss": 6,
               "extension": 7}

    def reserve_sms(self, n: int) -> None:
        """Leave n SMs out of K1's persistent grid (so a collective can run beside it); 0 = full grid."""
        _check(lib().mixfe_plan_reserve_sms(self._h, int(n)))

    def enable_timing(self, on: bool = True) -> None:
        _check(lib().mixfe_plan_enable_timing(self._h, 1 if on else 0))

    def kernel_time(self, which: str):
        """(total device ms, launches) of kernel family `which` since the last query (CUDA events)."""
        ms, n = ctypes.c_double(), ctypes.c_longlong()
        _check(lib().mixfe_plan_kernel_time(self._h, self.KERNELS[which], ctypes.byref(ms), ctypes.byref(n)))
        return ms.value, n.value

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            try:
                lib().mixfe_plan_destroy(self._h)
            finally:
                self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
