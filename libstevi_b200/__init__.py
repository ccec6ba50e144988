"""stevi_b200 — B200-native (sm_100a) dense-stereo hot path with LibStevi's ``StereoVision::Correlation`` surface.

The compute lives in ``libstevi_b200.so`` (hand-written CUDA behind the C ABI of ``include/stevi_b200/c_api.h``).
This package is the thin host-side mirror of the reference's call surface used by the tests and the benchmark;
C++ callers use the header facade under ``include/StereoVision``. There is no CPU fallback: importing
``libstevi_b200.api`` without the built library, or creating a context without a CUDA device, raises.
"""
from . import synth  # noqa: F401

__all__ = ["synth", "api"]
