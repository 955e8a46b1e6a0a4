"""dash_b200 -- B200-native evaluator for Dash's online garbled inference.

Only the hot path lives here: the CUDA kernels + C ABI (csrc/, built into
libdash_b200.so) and the ctypes mirror of the reference's GPU entry points
(api.py).  See DESIGN.md and INTEGRATION.md at the repository root.
"""
from .api import Circuit, DashB200Error, init, lib, nr_comps  # noqa: F401
