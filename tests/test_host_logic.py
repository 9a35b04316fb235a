"""CPU: host-side logic of the product package and the C-ABI surface (no compute calls)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest
import torch

import mixing_style_transfer_b200 as pkg
from mixing_style_transfer_b200 import _native
from mixing_style_transfer_b200.mixing_utils import hann_window_periodic, htk_mel_filterbank
from oracle import mixfe_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_and_exports_every_declared_symbol():
    path = pkg.build()
    assert os.path.exists(path)
    L = pkg.lib()
    header = open(os.path.join(ROOT, "include", "mixfe.h")).read()
    declared = set(re.findall(r"\b(mixfe_[a-z_]+)\s*\(", header))
    assert declared == set(_native.EXPORTS), declared ^ set(_native.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.mixfe_abi_version() == _native.ABI_VERSION == 2
    # mixfe_aug_params: gain, tilt gate, tilt sos, comp gate, comp threshold, comp ratio, bw gate, bw sos, reverb gate + pad
    assert ctypes.sizeof(_native.AugParams) == 4 * 4 + 4 * 4 + 4 * 6 * 8 + 4 * 4 + 4 * 4 + 4 * 4 + 4 * 4 + 4 * 12 * 8 + 8
    from mixing_style_transfer_b200.mixing_utils import _aug_dtype
    dt = _aug_dtype()                         # the numpy view draw_batch fills must match the C struct field by field
    assert dt.itemsize == ctypes.sizeof(_native.AugParams)
    for name, _ in _native.AugParams._fields_:
        assert dt.fields[name][1] == getattr(_native.AugParams, name).offset, name


def test_sass_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", _native.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out


def test_no_cpu_fallback_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    fe = pkg.MixingFeatureExtractor()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fe.extract_all_features(O.synthetic_clip(0, 4096))


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "mixing_style_transfer_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "/root/reference" not in src, f


@pytest.mark.parametrize("cfg", [(44100, 1024, 128), (44100, 2048, 80), (44100, 2048, 128), (22050, 1024, 64)])
def test_filterbank_and_window_match_oracle(cfg):
    sr, n_fft, n_mels = cfg
    assert torch.equal(htk_mel_filterbank(sr, n_fft, n_mels), O.htk_mel_filterbank(sr, n_fft, n_mels))
    assert torch.equal(hann_window_periodic(n_fft), O.hann_periodic(n_fft))


def test_feature_dim_and_names():
    fe = pkg.MixingFeatureExtractor()
    names = fe.feature_names()
    assert fe.get_feature_dim() == 64 and len(names) == 64
    assert [n.split('_')[0] for n in names] == [n.split('_')[0] for n in O.feature_layout()]
    fd = pkg.MixingFeatureExtractor(use_detailed_spectral=True, n_spectral_bins=32)
    assert fd.get_feature_dim() == 180 and len(fd.feature_names()) == 180
    assert (fe.sr, fe.n_fft, fe.hop_length, fe.n_mels) == (44100, 1024, 256, 128)


def test_butter_sos_matches_scipy():
    from scipy.signal import butter
    worst = 0.0
    for args in [(2, 2000, "high"), (2, 500, "low"), (4, 300.0, "high")] + [(4, fc, "low") for fc in np.linspace(4000, 11999.9, 17)]:
        a = butter(args[0], args[1], btype=args[2], fs=44100, output="sos")
        b = np.array(pkg.butter_sos(args[0], args[1], args[2], 44100))
        worst = max(worst, float(np.abs(a - b).max()))
    assert worst < 1e-13, worst


@pytest.mark.parametrize("seed,prob", [(1234, 1.0), (99, 0.5), (7, 0.5), (3, 0.0)])
def test_augmenter_draw_order_matches_reference_order(seed, prob):
    torch.manual_seed(seed)
    ref = O.draw_clip_aug(prob=prob)
    after_ref = torch.rand(1)
    torch.manual_seed(seed)
    got = pkg.AudioAugmenter(prob=prob).draw()
    after_got = torch.rand(1)
    assert torch.equal(after_ref, after_got)               # same number of draws consumed
    for n in O.STEMS:
        r = ref.stems[n]
        assert (got.gain_db[n] is None) == (r.gain_db is None)
        if r.gain_db is not None:
            assert torch.equal(got.gain_db[n], r.gain_db)
        assert got.tilt[n] == r.tilt and got.compress[n] == r.compress and got.cutoff_hz[n] == r.cutoff_hz
    assert (got.impulse is None) == (ref.impulse is None)
    if ref.impulse is not None:
        assert torch.equal(got.impulse, ref.impulse)


def test_aug_param_packing():
    torch.manual_seed(1234)
    aug = pkg.AudioAugmenter(prob=1.0)
    d = aug.draw()
    p = aug._pack([d])[0]
    assert p.reverb_on == 1 and all(p.tilt_on[i] == 1 and p.comp_on[i] == 1 and p.bw_on[i] == 1 for i in range(4))
    for i, n in enumerate(O.STEMS):
        assert p.gain_lin[i] == pytest.approx(float(10 ** (d.gain_db[n] / 20)), rel=1e-7)
        sos = O.butter_sos(4, d.cutoff_hz[n], "low", 44100)
        assert np.allclose(np.array(list(p.bw_sos[i])).reshape(2, 6), sos, rtol=0, atol=1e-14)


def test_mel_module_state_dict_keys():
    m = pkg.MelSpectrogramPreprocessor(44100, 1024, 256, 128)
    sd = m.state_dict()
    assert set(sd) == {"mel_transform.spectrogram.window", "mel_transform.mel_scale.fb"}
    assert sd["mel_transform.spectrogram.window"].shape == (1024,) and sd["mel_transform.mel_scale.fb"].shape == (513, 128)
    m.load_state_dict(sd, strict=True)


def test_host_fft_kernel_math():
    exe = "/tmp/mixfe_test_fft32"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "host", "test_fft32.cpp")])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout + out.stderr


def test_draw_batch_packs_what_per_clip_packing_packs():
    """draw_batch's vectorised decisions + numpy packing against the per-clip path (ClipAugmentation + _pack) fed with
    the same decisions: identical mixfe_aug_params, byte for byte (the vectorised butter_sos_batch included)."""
    import numpy as np
    aug = pkg.AudioAugmenter(prob=0.5)
    torch.manual_seed(7)
    batch = aug.draw_batch(16)
    assert len(batch) == 16 and batch.impulse is not None and batch.impulse.shape == (16, 22050)
    a = batch.arrays
    clips = []
    for b in range(16):
        d = pkg.ClipAugmentation()
        for s, n in enumerate(pkg.STEMS):
            d.gain_db[n] = torch.tensor([a["gain_db"][b, s]]) if a["gain_on"][b, s] else None
            d.tilt[n] = ("high" if a["tilt_high"][b, s] else "low") if a["tilt_on"][b, s] else None
            d.compress[n] = bool(a["comp_on"][b, s])
            d.cutoff_hz[n] = float(a["cutoff_hz"][b, s]) if a["bw_on"][b, s] else None
        d.impulse = batch.impulse[b] if a["reverb_on"][b] else None
        clips.append(d)
    ref = aug._pack(clips)
    assert ctypes.sizeof(batch.params) == ctypes.sizeof(ref)
    g = np.frombuffer(batch.params, dtype=_aug_dtype_for_test())
    w = np.frombuffer(ref, dtype=_aug_dtype_for_test())
    for name in g.dtype.names:
        if name == "_pad":
            continue
        if name == "gain_lin":                       # 10 ** (g / 20) on a vector vs on (1,) tensors: same fp32 kernel
            assert np.allclose(g[name], w[name], rtol=2e-7, atol=0), name
        elif name == "bw_sos":                       # numpy's vectorised complex arithmetic vs cmath: last-bit differences
            assert np.allclose(g[name], w[name], rtol=1e-13, atol=1e-16), name
        else:
            assert np.array_equal(g[name], w[name]), name
    # gates follow prob: about half of the 16 * 4 * 4 stem gates and of the 16 reverb gates fire
    fired = sum(int(a[k].sum()) for k in ("gain_on", "tilt_on", "comp_on", "bw_on"))
    assert 80 < fired < 176
    # a slice is a view of the same packed parameters
    sl = batch.slice(4, 9)
    assert len(sl) == 5 and sl.params[0].gain_lin[:] == batch.params[4].gain_lin[:]
    assert list(sl.reverb_rows) == [int(r) - 4 for r in batch.reverb_rows if 4 <= r < 9]
    # same seed, same decisions
    torch.manual_seed(7)
    again = aug.draw_batch(16)
    assert bytes(again.params) == bytes(batch.params) and torch.equal(again.impulse, batch.impulse)


def _aug_dtype_for_test():
    from mixing_style_transfer_b200.mixing_utils import _aug_dtype
    return _aug_dtype()


def test_butter_sos_batch_equals_scalar_design():
    import numpy as np
    from mixing_style_transfer_b200.mixing_utils import butter_sos_batch
    fcs = [4000.0, 5123.4, 8000.0, 11999.9]
    got = butter_sos_batch(4, fcs, "low", 44100)
    for i, fc in enumerate(fcs):
        assert np.allclose(got[i], np.asarray(pkg.butter_sos(4, fc, "low", 44100)), rtol=1e-13, atol=1e-16)
        assert np.allclose(got[i], O.butter_sos(4, fc, "low", 44100), rtol=1e-13, atol=1e-15)
    hp = butter_sos_batch(2, [2000.0], "high", 44100)
    assert np.allclose(hp[0], O.butter_sos(2, 2000, "high", 44100), rtol=1e-13, atol=1e-15)
