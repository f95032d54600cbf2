"""koboldcpp_b200 -- B200-native (sm_100a) implementation of koboldcpp's quantised-inference hot path.

The product is the C-ABI shared library `libkcpp_b200.so` (kernels + ggml_backend drop-in); this package is
the thin host-side Python mirror used by tests and bench.py.  There is no CPU fallback: importing `lib`
raises if the library is missing, and every compute call fails without a CUDA device.
"""
from .lib import B200Error, load  # noqa: F401
