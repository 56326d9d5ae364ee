"""kafka_inferenceengine_b200 -- B200-native drop-in for the per-pixel
assimilation path of KaFKA (QCDIS/KaFKA-InferenceEngine).

Mirrors the reference's package surface for that path (kafka/__init__.py:1-4):
``LinearKalman`` plus the ``inference`` functions, backed by hand-written
sm_100a CUDA kernels in ``csrc/`` behind the C ABI of ``include/kafka_b200.h``.
There is no CPU implementation in this package: without the built library (or
without a CUDA device) the compute entry points raise.
"""
from . import _lib                                         # noqa: F401
from .device import BlockDiag, DeviceVector, PixelIndex    # noqa: F401
from .inference import *                                   # noqa: F401,F403
from .input_output import *                                # noqa: F401,F403
from .linear_kf import LinearKalman                        # noqa: F401
from .observation_operators import *                       # noqa: F401,F403

__version__ = "0.1.0"
