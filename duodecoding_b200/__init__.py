"""duodecoding_b200 - B200-native speculative-decoding verification path.

Drop-in for the verification hot path of Nanji-Huaji/DuoDecoding
(src/utils.py, src/decoding_ops.py, src/decoding_types.py,
src/proposal_utils.py, src/communication.py compression helpers,
src/model_gpu.py): same function names and signatures, hand-written sm_100a
CUDA kernels behind a C ABI (include/duodec.h).  Importing the package loads
``libduodec.so``; there is no CPU or eager fallback.
"""

from . import _lib, ops  # noqa: F401  (loads libduodec.so, fails loudly if it is missing)

__version__ = "0.1.0"
__all__ = ["ops", "_lib"]
