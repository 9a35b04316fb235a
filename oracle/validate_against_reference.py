"""Pin the oracle against the real reference (build container only; TEST INFRASTRUCTURE).

Imports the reference's own modules read-only from /root/reference/src and compares every oracle
function with them on shared inputs.  Run:  python oracle/validate_against_reference.py
The GPU box has no /root/reference; there the committed goldens (tests/golden/) stand in.
"""
import os
import sys
import warnings

os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
sys.dont_write_bytecode = True
warnings.filterwarnings("ignore")
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

import numpy as np
import torch

from oracle import mixfe_oracle as O

REF_SRC = "/root/reference/src"


def load_reference():
    if not os.path.isdir(REF_SRC):
        raise SystemExit("reference checkout not present; nothing to validate against")
    sys.path.insert(0, REF_SRC)
    import mixing_utils                                   # noqa: E402
    from model import MelSpectrogramPreprocessor          # noqa: E402
    return mixing_utils, MelSpectrogramPreprocessor


def maxdiff(a, b):
    return float((a.double() - b.double()).abs().max())


def main():
    mu, RefMel = load_reference()
    ok = True

    def check(name, cond, detail=""):
        nonlocal ok
        ok &= bool(cond)
        print(("PASS " if cond else "FAIL ") + name + (" -- " + detail if detail else ""))

    for (n_fft, hop, n_mels) in [(1024, 256, 128), (2048, 512, 80), (2048, 512, 128)]:
        cfg = O.FrontEndConfig(44100, n_fft, hop, n_mels)
        ref_fe = mu.MixingFeatureExtractor(44100, n_fft, hop, n_mels)
        fe = O.MelFrontEnd(cfg)
        check(f"fb bit-exact {n_fft}/{hop}/{n_mels}", torch.equal(fe.fb, ref_fe.mel_transform.mel_scale.fb))
        check(f"window bit-exact {n_fft}", torch.equal(fe.window, ref_fe.mel_transform.spectrogram.window))
        for T in (n_fft // 2 + 1, 4097, 44100):
            clip = O.synthetic_clip(7, T)
            m_ref = ref_fe.mel_transform(clip["drums"])
            m_or = fe.mel_power(clip["drums"])
            check(f"mel power bit-exact T={T}", torch.equal(m_ref, m_or), f"maxdiff {maxdiff(m_ref, m_or):.3g}")
        for mk, T in ((O.synthetic_clip, 66150), (O.musical_clip, 66150), (O.synthetic_clip, 441000)):
            clip = mk(1234, T)
            f_ref = ref_fe.extract_all_features(clip)
            f_or = O.mixing_features(clip, cfg, fe=fe)
            check(f"features {mk.__name__} T={T}", torch.equal(f_ref, f_or), f"maxdiff {maxdiff(f_ref, f_or):.3g}")
        batch = {n: torch.stack([O.musical_clip(s, 30000)[n] for s in (1, 2)]) for n in O.STEMS}
        lm_ref = RefMel(44100, n_fft, hop, n_mels)(batch)
        lm_or = O.log_mel(batch, cfg)
        check("log-mel bit-exact", torch.equal(lm_ref, lm_or), f"maxdiff {maxdiff(lm_ref, lm_or):.3g}")

    # detailed-spectral mode
    for nb in (32, 128, 16):
        cfg = O.FrontEndConfig(use_detailed_spectral=True, n_spectral_bins=nb)
        ref_fe = mu.MixingFeatureExtractor(use_detailed_spectral=True, n_spectral_bins=nb)
        clip = O.musical_clip(5, 50000)
        f_ref, f_or = ref_fe.extract_all_features(clip), O.mixing_features(clip, cfg)
        check(f"detailed spectral n_bins={nb} dim={cfg.feature_dim}",
              f_ref.shape[0] == cfg.feature_dim == ref_fe.get_feature_dim() and torch.equal(f_ref, f_or),
              f"maxdiff {maxdiff(f_ref, f_or):.3g}")

    # special inputs: silence, L=R, L=-R, only vocals active
    ref_fe = mu.MixingFeatureExtractor()
    T = 44100
    z = {n: torch.zeros(2, T) for n in O.STEMS}
    check("silence", torch.equal(ref_fe.extract_all_features(z), O.mixing_features(z)))
    g = torch.Generator().manual_seed(1)
    m = torch.randn(T, generator=g) * 0.1
    for label, pair in (("L=R", torch.stack([m, m])), ("L=-R", torch.stack([m, -m]))):
        c = {n: pair.clone() for n in O.STEMS}
        check(label, torch.equal(ref_fe.extract_all_features(c), O.mixing_features(c)))
    c = {n: torch.zeros(2, T) for n in O.STEMS}
    c["vocals"] = torch.stack([m, 0.5 * m])
    check("only vocals", torch.equal(ref_fe.extract_all_features(c), O.mixing_features(c)))
    try:
        O.mixing_features({n: torch.zeros(2, 400) for n in O.STEMS})
        check("T<=n_fft/2 raises", False)
    except RuntimeError:
        check("T<=n_fft/2 raises", True)

    # filters
    from scipy.signal import butter, sosfilt
    worst = 0.0
    for args in [(2, 2000, "high"), (2, 500, "low")] + [(4, fc, "low") for fc in np.linspace(4000, 11999.9, 23)]:
        a = butter(args[0], args[1], btype=args[2], fs=44100, output="sos")
        b = O.butter_sos(args[0], args[1], args[2], 44100)
        worst = max(worst, float(np.abs(a - b).max()))
    check("butter_sos vs scipy", worst < 1e-13, f"max coeff diff {worst:.3g}")
    x = O.synthetic_clip(3, 20000)["bass"].numpy()
    for sos in (butter(2, 500, btype="low", fs=44100, output="sos"), butter(4, 7000.5, btype="low", fs=44100, output="sos")):
        d = float(np.abs(sosfilt(sos, x, axis=-1) - O.sosfilt_df2t(sos, x)).max())
        check("sosfilt_df2t vs scipy.sosfilt", d < 1e-15, f"maxdiff {d:.3g}")

    # reverb forms
    h = (torch.randn(22050, generator=g) * 0.1).numpy()
    xs = x[0, :6000]
    d_fft = float(np.abs(O.xcorr_same_direct(xs, h) - O.xcorr_same_fft(xs, h)).max())
    check("xcorr fft vs direct", d_fft < 1e-6, f"maxdiff {d_fft:.3g}")

    # whole augmentation chain: same seed -> same decisions -> same waveforms
    aug_ref = mu.AudioAugmenter(prob=1.0)
    for seed, prob, T in ((1234, 1.0, 30000), (99, 0.5, 30000), (7, 0.5, 30000), (11, 0.5, 30000)):
        clip = O.musical_clip(seed, T)
        aug_ref.prob = prob
        torch.manual_seed(seed)
        y_ref = aug_ref.augment_stems(clip)
        torch.manual_seed(seed)
        params = O.draw_clip_aug(prob=prob)
        y_dir = O.apply_clip_aug(clip, params, fft_reverb=False)
        y_fft = O.apply_clip_aug(clip, params, fft_reverb=True)
        scale = max(float(v.abs().max()) for v in y_ref.values())
        d1 = max(maxdiff(y_ref[n], y_dir[n]) for n in O.STEMS) / scale
        d2 = max(maxdiff(y_ref[n], y_fft[n]) for n in O.STEMS) / scale
        check(f"augment_stems seed={seed} prob={prob} reverb={'on' if params.impulse is not None else 'off'}",
              # the reference's 22 050-tap fp32 conv1d carries ~5e-6 of its own accumulation noise
              d1 < 2e-5 and d2 < 2e-5, f"rel maxdiff direct {d1:.3g} fft {d2:.3g}")
    print("ALL PASS" if ok else "SOME FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
