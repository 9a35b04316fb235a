"""mixing_style_transfer_b200 -- B200-native (sm_100a) DSP front end of barry-mir/mixing-style-transfer.

Public surface = the reference's three hot-path classes, same names and signatures:
``MixingFeatureExtractor``, ``AudioAugmenter`` (reference: src/mixing_utils.py) and
``MelSpectrogramPreprocessor`` (reference: src/model.py:17-67), all running on the hand-written
CUDA kernels of ``csrc/`` through the C ABI declared in ``include/mixfe.h``.
"""
from ._native import STEMS, Plan, build, launch_count, lib
from .mixing_utils import AudioAugmenter, BatchAugmentation, ClipAugmentation, MixingFeatureExtractor, butter_sos
from .model import MelSpectrogramPreprocessor
from .loss import MultiResolutionSTFTLoss

__all__ = ["STEMS", "Plan", "build", "lib", "launch_count", "MixingFeatureExtractor", "AudioAugmenter",
           "ClipAugmentation", "BatchAugmentation", "MelSpectrogramPreprocessor", "MultiResolutionSTFTLoss", "butter_sos"]
