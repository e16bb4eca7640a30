"""llamacog_b200 -- B200-native (sm_100a) kernels and ggml backend plugin for llamacog's quantized
transformer forward pass.  Python here is only the host-side mirror used by tests and bench.py:
ctypes bindings of the C ABI in include/b200_ops.h.  There is NO CPU fallback: importing `ops`
without the built CUDA library raises."""
from . import build  # noqa: F401

__all__ = ["build", "ops"]
