"""sph_b200 — B200-native back-end for the SPH (Superpixel Hierarchies) hot path.

Only what the path needs: csrc/ (sm_100a kernels + the C ABI of include/sphb.h),
capi.py (ctypes view of that ABI), host.py (Python mirror of the reference's
stage interface, for tests/bench) and synth.py (the synthetic image stacks of
BASELINE.json).  No CPU fallback: without the built library and a B200 the
compute calls raise.
"""
from . import capi  # noqa: F401

__all__ = ["capi"]
