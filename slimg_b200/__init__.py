"""slimg_b200 -- B200-native implementation of slimg's per-image pixel pipeline.

The public surface mirrors the reference's Python package (bindings/python/slimg/__init__.py:3-19):
``Image``, ``Resize``, ``Crop``, ``Extend``, ``resize``, ``crop``, ``extend``, ``convert`` ...; the
work is done by hand-written sm_100a CUDA kernels behind the C ABI of libslimg_b200.so
(include/slimg_b200.h).  There is no CPU fallback.
"""
from ._api import (  # noqa: F401
    Format,
    Image,
    Result,
    Resize,
    Crop,
    Extend,
    SlimgError,
    JpegCoefficients,
    convert,
    convert_pixels,
    crop_image as crop,
    extend_image as extend,
    resize_image as resize,
    resize_crop_extend,
    jpeg_frontend,
    jpeg_quant_tables,
    decode,
    open,
    optimize,
    optimize_file,
)

__all__ = [
    "Format", "Image", "Result", "Resize", "Crop", "Extend", "SlimgError", "JpegCoefficients",
    "open", "decode", "convert", "convert_pixels", "crop", "extend", "resize", "resize_crop_extend",
    "jpeg_frontend", "jpeg_quant_tables", "optimize", "optimize_file",
]
