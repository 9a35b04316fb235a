"""Generate tests/golden/*.npz from the UNMODIFIED reference (build container only).

Imports /root/reference/src/{mixing_utils,model}.py read-only and records its outputs on inputs
that tests regenerate from seeds (oracle.mixfe_oracle.synthetic_clip / musical_clip), so only
outputs are stored.  Run:  python tests/golden/make_golden.py
"""
import os
import sys
import warnings

os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
sys.dont_write_bytecode = True
warnings.filterwarnings("ignore")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/src")

import numpy as np
import torch

import mixing_utils as ref_mu                       # the reference itself
from model import MelSpectrogramPreprocessor as RefMel
from oracle import mixfe_oracle as O


def main():
    out = {}
    fe = ref_mu.MixingFeatureExtractor()
    for mk, seed, T in ((O.synthetic_clip, 1234, 66150), (O.musical_clip, 1234, 66150),
                        (O.synthetic_clip, 1234, 441000), (O.musical_clip, 9, 4097), (O.musical_clip, 21, 2048)):
        out[f"features/{mk.__name__}/{seed}/{T}"] = fe.extract_all_features(mk(seed, T)).numpy()
    for nb in (32, 16, 128):
        fd = ref_mu.MixingFeatureExtractor(use_detailed_spectral=True, n_spectral_bins=nb)
        out[f"features_detailed{nb}/musical_clip/5/50000"] = fd.extract_all_features(O.musical_clip(5, 50000)).numpy()
    # special inputs
    T = 44100
    z = {n: torch.zeros(2, T) for n in O.STEMS}
    out["features/silence"] = fe.extract_all_features(z).numpy()
    g = torch.Generator().manual_seed(1)
    m = torch.randn(T, generator=g) * 0.1
    out["features/L_eq_R"] = fe.extract_all_features({n: torch.stack([m, m]) for n in O.STEMS}).numpy()
    out["features/L_eq_negR"] = fe.extract_all_features({n: torch.stack([m, -m]) for n in O.STEMS}).numpy()
    c = {n: torch.zeros(2, T) for n in O.STEMS}
    c["vocals"] = torch.stack([m, 0.5 * m])
    out["features/only_vocals"] = fe.extract_all_features(c).numpy()
    # log-mel (batched), subsampled
    batch = {n: torch.stack([O.musical_clip(s, 30000)[n] for s in (1, 2)]) for n in O.STEMS}
    lm = RefMel(44100, 1024, 256, 128)(batch)
    out["logmel/musical_clip/1,2/30000/sub"] = lm[:, :, ::4, ::8].numpy()
    out["logmel/musical_clip/1,2/30000/stats"] = np.array([lm.mean(), lm.min(), lm.max()], dtype=np.float64)
    lm = RefMel(44100, 1024, 256, 128)({n: v[None] for n, v in O.synthetic_clip(7, 8192).items()})
    out["logmel/synthetic_clip/7/8192/full"] = lm.numpy()
    # other STFT configurations the tree uses
    for (n_fft, hop, n_mels) in ((2048, 512, 80), (2048, 512, 128)):
        lm = RefMel(44100, n_fft, hop, n_mels)({n: v[None] for n, v in O.musical_clip(4, 20000).items()})
        out[f"logmel_{n_fft}_{hop}_{n_mels}/musical_clip/4/20000/sub"] = lm[:, :, ::2, ::4].numpy()
        f2 = ref_mu.MixingFeatureExtractor(44100, n_fft, hop, n_mels).extract_all_features(O.musical_clip(4, 20000))
        out[f"features_{n_fft}_{hop}_{n_mels}/musical_clip/4/20000"] = f2.numpy()
    # augmentation: same seed -> same decisions -> same waveforms
    aug = ref_mu.AudioAugmenter(prob=1.0)
    for seed, prob, T in ((1234, 1.0, 30000), (99, 0.5, 30000), (7, 0.5, 30000), (11, 0.5, 30000)):
        clip = O.musical_clip(seed, T)
        aug.prob = prob
        torch.manual_seed(seed)
        y = aug.augment_stems(clip)
        for n in O.STEMS:
            out[f"augment/{seed}/{prob}/{T}/{n}/sub5"] = y[n][:, ::5].numpy()
            out[f"augment/{seed}/{prob}/{T}/{n}/energy"] = np.array([float((y[n].double() ** 2).mean())])
    # song_A known-answer vector (input = the reference's own asset, not shipped)
    wav = "/root/reference/assets/song_A.wav"
    if os.path.exists(wav):
        from scipy.io import wavfile
        sr, x = wavfile.read(wav)
        x = torch.from_numpy(x.astype(np.float32) / 32768.0).t().contiguous()
        T = 441000
        clip = {n: x[:, o:o + T].contiguous() for n, o in zip(O.STEMS, (0, 66150, 132300, 198450))}
        out["features/song_A"] = fe.extract_all_features(clip).numpy()
        lm = RefMel(44100, 1024, 256, 128)({n: v[None] for n, v in clip.items()})
        out["logmel/song_A/stats"] = np.array([lm.mean(), lm.min(), lm.max(), lm[0, 0, 64, 861], lm[0, 7, 127, 1722]],
                                              dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "reference_outputs.npz"), **out)
    print("wrote", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "reference_outputs.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
