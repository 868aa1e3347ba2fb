"""slimg_b200 -- B200-native implementation of slimg's per-image pixel pipeline.

``slimg_b200._lowlevel`` is a drop-in for the UniFFI-generated ``slimg._lowlevel`` module of the reference's Python
binding: the reference's own wrapper (bindings/python/slimg/_types.py) runs on it UNMODIFIED once
``slimg_b200.install_as_slimg_lowlevel()`` has registered it under that name.  The work is done by hand-written
sm_100a CUDA kernels behind the C ABI of libslimg_b200.so (include/slimg_b200.h).  There is no CPU fallback.
"""
from ._lowlevel import install_as_slimg_lowlevel  # noqa: F401

__all__ = ["install_as_slimg_lowlevel"]
