"""signalizer_b200 — B200-native (sm_100a) implementation of Signalizer's analysis hot path.

Host-side Python mirror of the reference's TransformConstant/TransformPair interface over the
C ABI in include/sgz.h (libsgz_cuda.so).  There is no CPU fallback: importing the compute
classes without the built CUDA library raises."""
from . import _abi  # noqa: F401
from ._abi import default_config, copy_config  # noqa: F401

__all__ = ["_abi", "default_config", "copy_config"]
