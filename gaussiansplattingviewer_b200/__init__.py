"""B200-native forward Gaussian-splat renderer behind the GaussianSplattingViewer renderer interface.

Host side (this package) mirrors the reference's Python surface; all rendering work happens in
hand-written sm_100a CUDA kernels inside ``csrc/libgsr.so``, reached through the C ABI declared in
``include/gsr.h``.  There is no CPU fallback: constructing a renderer without the CUDA library
raises.
"""
from . import util, util_gau, scenes  # noqa: F401
