"""audiossl_b200 -- B200-native drop-in for the audio-ssl upstream input path.

Exports the reference's names for this path (log-mel front end, two-view augmentation) backed by
hand-written sm_100a CUDA kernels behind the C ABI in ``include/assl.h``.  There is no CPU fallback.

    from audiossl_b200 import MelSpectrogramLibrosa, extract_log_mel_spectrogram, AugmentationModule
"""
from . import _native
from .augmentations import (Kmix, MixupBYOLA, NormalizeBatch, PrecomputedNorm, RandomResizeCrop,  # noqa: F401
                            RunningNorm)
from .datamodule import BaselineDataModule, DeviceBatchLoader, PackedClips, WaveformDataset  # noqa: F401
from .engine import StatsExchange, ViewEngine, shard_bounds  # noqa: F401
from .frontend import (HostPipeline, MelSpectrogramLibrosa, Plan, extract_log_mel_spectrogram, extract_window,  # noqa: F401
                       extract_windows, extract_windows_packed)
from .module import AugmentationModule, UpstreamInputPipeline  # noqa: F401
from .sampler import FastSampler, ReferenceSampler, ViewParams  # noqa: F401

build = _native.build
__version__ = "0.1.0"
