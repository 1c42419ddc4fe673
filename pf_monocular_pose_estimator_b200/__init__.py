"""pf_monocular_pose_estimator_b200 -- B200-native particle-filter tracking core of PF-MPE.

The product is libpfmpe.so (hand-written sm_100a CUDA behind the C ABI of include/pfmpe.h);
`capi` is its ctypes mirror, `scenes` the synthetic-scene generator used by tests and bench.
Nothing in this package computes the hot path on the CPU.
"""
from . import capi, scenes  # noqa: F401

__all__ = ["capi", "scenes"]
