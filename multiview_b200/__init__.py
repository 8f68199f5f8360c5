"""multiview_b200 -- B200 (sm_100a) implementation of multiview's per-frame stereo reconstruction hot path.

Compute lives in ``libmvstereo.so`` (hand-written CUDA behind the C ABI of ``include/mvstereo.h``);
``multiview_b200.api`` is the host-side mirror of the reference's operator interface for that path.
"""
__all__ = ["api", "synth", "build"]
