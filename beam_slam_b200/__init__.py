"""beam_slam_b200 — B200-native Levenberg-Marquardt back end behind fuse_core::Graph::optimize.

Only what the hot path needs: csrc/ (CUDA kernels + C-ABI, built to lib/libbsnls.so), host/ (C++ mirror of the
fuse graph interface + flattener), and thin ctypes plumbing. No CPU solve path exists in this package.
"""
from . import cabi  # noqa: F401
