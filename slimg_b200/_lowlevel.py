"""ctypes binding of libslimg_b200.so (include/slimg_b200.h).

Plays the part of the UniFFI-generated ``slimg._lowlevel`` module of the reference's Python
binding (bindings/python/slimg/_types.py imports exactly these names): ``Format``,
``ImageData``, ``ResizeMode.WIDTH(value=)`` ..., ``CropMode``, ``ExtendMode``, ``FillColor``,
``PipelineOptions``, the functions ``resize / crop / extend / convert ...`` and the ``SlimgError``
exception namespace.  Pixel work is done by the CUDA library; there is no CPU fallback -- if the
library is missing or no CUDA device is usable the call raises.
"""
from __future__ import annotations

import ctypes as C
import enum
import os
from dataclasses import dataclass
from typing import Optional

from . import _build

# --------------------------------------------------------------------------------------------
# C structs (include/slimg_b200.h)
# --------------------------------------------------------------------------------------------


class CImage(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("data", C.c_void_p), ("len", C.c_size_t)]


class CResizeMode(C.Structure):
    _fields_ = [("tag", C.c_int32), ("a", C.c_uint32), ("b", C.c_uint32), ("factor", C.c_double)]


class CCropMode(C.Structure):
    _fields_ = [("tag", C.c_int32), ("x", C.c_uint32), ("y", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32)]


class CExtendMode(C.Structure):
    _fields_ = [("tag", C.c_int32), ("width", C.c_uint32), ("height", C.c_uint32)]


class CFillColor(C.Structure):
    _fields_ = [("tag", C.c_int32), ("r", C.c_uint8), ("g", C.c_uint8), ("b", C.c_uint8), ("a", C.c_uint8)]


class CPipelineOptions(C.Structure):
    _fields_ = [("has_resize", C.c_int32), ("resize", CResizeMode),
                ("has_crop", C.c_int32), ("crop", CCropMode),
                ("has_extend", C.c_int32), ("extend", CExtendMode),
                ("has_fill", C.c_int32), ("fill", CFillColor)]


class CJpegGeometry(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("width", "height", "y_wb", "y_hb", "y_WB", "y_HB", "c_wb", "c_hb", "c_WB", "c_HB")]


class CJpegCoefs(C.Structure):
    _fields_ = [("geom", CJpegGeometry), ("qt_luma", C.c_uint16 * 64), ("qt_chroma", C.c_uint16 * 64),
                ("y", C.POINTER(C.c_int16)), ("cb", C.POINTER(C.c_int16)), ("cr", C.POINTER(C.c_int16)),
                ("y_uq", C.POINTER(C.c_int16)), ("cb_uq", C.POINTER(C.c_int16)), ("cr_uq", C.POINTER(C.c_int16))]


class CPlanes(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("kind", C.c_uint32), ("chroma_width", C.c_uint32),
                ("chroma_height", C.c_uint32), ("bytes_per_sample", C.c_uint32), ("has_alpha", C.c_int32),
                ("y", C.c_void_p), ("cb", C.c_void_p), ("cr", C.c_void_p), ("a", C.c_void_p)]


class CBatchConfig(C.Structure):
    _fields_ = [("n_devices", C.c_int32), ("streams_per_device", C.c_int32), ("device_ids", C.POINTER(C.c_int32))]


ORDER_CONVERT, ORDER_CHAIN = 0, 1
QT_ANNEX_K, QT_MOZJPEG = 0, 1
JPEG_UNQUANTISED, JPEG_DERINGING = 1, 2
PLANES_WEBP_YUV420, PLANES_AVIF_YCBCR444_8, PLANES_AVIF_YCBCR444_10 = 1, 2, 3
PATH_NONE, PATH_REMAP, PATH_GENERIC, PATH_FEWTAPS, PATH_HFIRST, PATH_VFIRST, PATH_TC, PATH_TILE = range(8)
PATH_NAMES = ("none", "remap", "generic", "fewtaps", "hfirst", "vfirst", "tc", "tile")

# every symbol include/slimg_b200.h declares (tests/test_abi.py checks the .so exports them all)
ABI_SYMBOLS = (
    "slimg_b200_calculate_dimensions", "slimg_b200_resize_output_dims", "slimg_b200_calculate_crop_region",
    "slimg_b200_calculate_extend_region", "slimg_b200_pipeline_output_dims", "slimg_b200_crop",
    "slimg_b200_extend", "slimg_b200_resize", "slimg_b200_pipeline", "slimg_b200_image_free",
    "slimg_b200_jpeg_geometry", "slimg_b200_jpeg_quant_table", "slimg_b200_jpeg_frontend",
    "slimg_b200_jpeg_coefs_free", "slimg_b200_pipeline_batch_device", "slimg_b200_jpeg_frontend_batch_device",
    "slimg_b200_pipeline_batch_host", "slimg_b200_last_error", "slimg_b200_abi_version",
    "slimg_b200_device_count", "slimg_b200_kernel_launches", "slimg_b200_host_alloc", "slimg_b200_host_free",
    "slimg_b200_pipeline_source_window", "slimg_b200_jpeg_frontend_ex", "slimg_b200_pipeline_jpeg",
    "slimg_b200_jpeg_frontend_batch_device_ex", "slimg_b200_pipeline_jpeg_batch_device",
    "slimg_b200_pipeline_jpeg_batch_host", "slimg_b200_last_resize_path", "slimg_b200_path_launches",
    "slimg_b200_pipeline_planes", "slimg_b200_planes_free", "slimg_b200_color_planes_batch_device",
    "slimg_b200_webp_planes_batch_device_static",
)

_lib: Optional[C.CDLL] = None


def library() -> C.CDLL:
    """Load libslimg_b200.so (building it in-tree first when the sources are newer)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    override = os.environ.get("SLIMG_B200_LIB")         # A/B of two builds of the same sources
    if override:
        path = override
    elif _build.needs_build():
        path = _build.build()
    if not os.path.exists(path):
        raise ImportError(f"libslimg_b200.so not found at {path}; run `python -m slimg_b200._build`")
    lib = C.CDLL(path)
    u32p, vp = C.POINTER(C.c_uint32), C.c_void_p
    sigs = {
        "slimg_b200_calculate_dimensions": (C.c_int, [C.c_uint32, C.c_uint32, C.POINTER(CResizeMode), u32p, u32p]),
        "slimg_b200_resize_output_dims": (C.c_int, [C.c_uint32, C.c_uint32, C.POINTER(CResizeMode), u32p, u32p]),
        "slimg_b200_calculate_crop_region": (C.c_int, [C.c_uint32, C.c_uint32, C.POINTER(CCropMode), u32p, u32p, u32p, u32p]),
        "slimg_b200_calculate_extend_region": (C.c_int, [C.c_uint32, C.c_uint32, C.POINTER(CExtendMode), u32p, u32p, u32p, u32p]),
        "slimg_b200_pipeline_output_dims": (C.c_int, [C.c_uint32, C.c_uint32, C.POINTER(CPipelineOptions), C.c_int, u32p, u32p]),
        "slimg_b200_pipeline_source_window": (C.c_int, [C.c_uint32, C.c_uint32, C.POINTER(CPipelineOptions), C.c_int, u32p, u32p, u32p, u32p]),
        "slimg_b200_crop": (C.c_int, [C.POINTER(CImage), C.POINTER(CCropMode), C.POINTER(CImage)]),
        "slimg_b200_extend": (C.c_int, [C.POINTER(CImage), C.POINTER(CExtendMode), C.POINTER(CFillColor), C.POINTER(CImage)]),
        "slimg_b200_resize": (C.c_int, [C.POINTER(CImage), C.POINTER(CResizeMode), C.POINTER(CImage)]),
        "slimg_b200_pipeline": (C.c_int, [C.POINTER(CImage), C.POINTER(CPipelineOptions), C.c_int, C.POINTER(CImage)]),
        "slimg_b200_image_free": (None, [C.POINTER(CImage)]),
        "slimg_b200_jpeg_geometry": (None, [C.c_uint32, C.c_uint32, C.POINTER(CJpegGeometry)]),
        "slimg_b200_jpeg_quant_table": (None, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint16)]),
        "slimg_b200_jpeg_frontend": (C.c_int, [C.POINTER(CImage), C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.POINTER(CJpegCoefs)]),
        "slimg_b200_jpeg_coefs_free": (None, [C.POINTER(CJpegCoefs)]),
        "slimg_b200_pipeline_batch_device": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(CPipelineOptions), C.c_int, vp, C.c_size_t, vp]),
        "slimg_b200_jpeg_frontend_batch_device": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), vp, vp, vp, vp]),
        "slimg_b200_pipeline_batch_host": (C.c_int32, [C.POINTER(CImage), C.c_uint32, C.POINTER(CPipelineOptions), C.c_int, C.POINTER(CBatchConfig), C.POINTER(CImage), C.POINTER(C.c_int)]),
        "slimg_b200_last_error": (C.c_char_p, []),
        "slimg_b200_abi_version": (C.c_int32, []),
        "slimg_b200_device_count": (C.c_int32, []),
        "slimg_b200_kernel_launches": (C.c_uint64, []),
        "slimg_b200_host_alloc": (vp, [C.c_size_t]),
        "slimg_b200_host_free": (None, [vp]),
        "slimg_b200_jpeg_frontend_ex": (C.c_int, [C.POINTER(CImage), C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.c_uint32, C.POINTER(CJpegCoefs)]),
        "slimg_b200_pipeline_jpeg": (C.c_int, [C.POINTER(CImage), C.POINTER(CPipelineOptions), C.c_int, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.c_uint32, C.POINTER(CJpegCoefs)]),
        "slimg_b200_jpeg_frontend_batch_device_ex": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), vp, vp, vp, vp, vp, vp, C.c_uint32, vp]),
        "slimg_b200_pipeline_jpeg_batch_device": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(CPipelineOptions), C.c_int, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), vp, C.c_size_t, vp, vp, vp, vp, vp, vp, C.c_uint32, vp]),
        "slimg_b200_pipeline_jpeg_batch_host": (C.c_int32, [C.POINTER(CImage), C.c_uint32, C.POINTER(CPipelineOptions), C.c_int, C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.c_uint32, C.POINTER(CBatchConfig), C.POINTER(CJpegCoefs), C.POINTER(C.c_int)]),
        "slimg_b200_pipeline_planes": (C.c_int, [C.POINTER(CImage), C.POINTER(CPipelineOptions), C.c_int, C.c_int, C.POINTER(CPlanes)]),
        "slimg_b200_planes_free": (None, [C.POINTER(CPlanes)]),
        "slimg_b200_color_planes_batch_device": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, vp, vp, vp, vp, vp, vp]),
        "slimg_b200_webp_planes_batch_device_static": (C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp, vp]),
        "slimg_b200_last_resize_path": (C.c_int32, []),
        "slimg_b200_path_launches": (C.c_uint64, [C.c_int32]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


# --------------------------------------------------------------------------------------------
# Errors: the ten SlimgError variants (crates/slimg-ffi/src/lib.rs:195-227)
# --------------------------------------------------------------------------------------------


class SlimgError(Exception):
    """Base class and namespace of the error variants (``SlimgError.Crop`` ...)."""


def _variant(name: str):
    cls = type(name, (SlimgError,), {"__doc__": f"SlimgError::{name}"})
    setattr(SlimgError, name, cls)
    return cls


_STATUS_TO_ERROR = {
    1: _variant("UnsupportedFormat"), 2: _variant("UnknownFormat"), 3: _variant("EncodingNotSupported"),
    4: _variant("Decode"), 5: _variant("Encode"), 6: _variant("Resize"), 7: _variant("Crop"),
    8: _variant("Extend"), 9: _variant("Io"), 10: _variant("Image"),
}


def _check(status: int) -> None:
    if status:
        msg = library().slimg_b200_last_error().decode("utf-8", "replace")
        raise _STATUS_TO_ERROR.get(status, SlimgError)(msg)


# --------------------------------------------------------------------------------------------
# Value types with the UniFFI-generated shape
# --------------------------------------------------------------------------------------------


class Format(enum.Enum):
    JPEG = 1
    PNG = 2
    WEB_P = 3
    AVIF = 4
    JXL = 5
    QOI = 6


@dataclass
class ImageData:
    width: int
    height: int
    data: bytes


class ResizeMode:
    """``ResizeMode.WIDTH(value=800)`` etc. (crates/slimg-ffi/src/lib.rs:40-53)."""

    def __init__(self, tag, a=0, b=0, factor=0.0):
        self._c = CResizeMode(tag, a, b, factor)

    def __repr__(self):
        c = self._c
        return f"ResizeMode(tag={c.tag}, a={c.a}, b={c.b}, factor={c.factor})"

    @classmethod
    def WIDTH(cls, value): return cls(1, _u32(value, "value"))
    @classmethod
    def HEIGHT(cls, value): return cls(2, _u32(value, "value"))
    @classmethod
    def EXACT(cls, width, height): return cls(3, _u32(width, "width"), _u32(height, "height"))
    @classmethod
    def FIT(cls, max_width, max_height): return cls(4, _u32(max_width, "max_width"), _u32(max_height, "max_height"))
    @classmethod
    def SCALE(cls, factor): return cls(5, 0, 0, float(factor))


class CropMode:
    def __init__(self, tag, x=0, y=0, width=0, height=0):
        self._c = CCropMode(tag, x, y, width, height)

    @classmethod
    def REGION(cls, x, y, width, height):
        return cls(1, _u32(x, "x"), _u32(y, "y"), _u32(width, "width"), _u32(height, "height"))

    @classmethod
    def ASPECT_RATIO(cls, width, height): return cls(2, 0, 0, _u32(width, "width"), _u32(height, "height"))


class ExtendMode:
    def __init__(self, tag, width, height):
        self._c = CExtendMode(tag, width, height)

    @classmethod
    def ASPECT_RATIO(cls, width, height): return cls(1, _u32(width, "width"), _u32(height, "height"))
    @classmethod
    def SIZE(cls, width, height): return cls(2, _u32(width, "width"), _u32(height, "height"))


class FillColor:
    def __init__(self, tag, r=0, g=0, b=0, a=0):
        self._c = CFillColor(tag, r, g, b, a)

    @classmethod
    def SOLID(cls, r, g, b, a): return cls(1, _u8(r), _u8(g), _u8(b), _u8(a))
    @classmethod
    def TRANSPARENT(cls): return cls(2)


@dataclass
class PipelineOptions:
    format: Format
    quality: int
    resize: Optional[ResizeMode] = None
    crop: Optional[CropMode] = None
    extend: Optional[ExtendMode] = None
    fill_color: Optional[FillColor] = None

    def _to_c(self) -> CPipelineOptions:
        return make_options(self.resize, self.crop, self.extend, self.fill_color)


@dataclass
class PipelineResult:
    data: bytes
    format: Format
    width: int
    height: int


def _u32(v, name) -> int:
    if not isinstance(v, int) or isinstance(v, bool) or not 0 <= v <= 0xFFFFFFFF:
        raise ValueError(f"{name} must be an integer in 0..2^32-1, got {v!r}")
    return v


def _u8(v) -> int:
    if not isinstance(v, int) or not 0 <= v <= 255:
        raise ValueError(f"colour channel must be an integer in 0-255, got {v!r}")
    return v


def make_options(resize=None, crop=None, extend=None, fill_color=None) -> CPipelineOptions:
    o = CPipelineOptions()
    if resize is not None: o.has_resize, o.resize = 1, resize._c
    if crop is not None: o.has_crop, o.crop = 1, crop._c
    if extend is not None: o.has_extend, o.extend = 1, extend._c
    if fill_color is not None: o.has_fill, o.fill = 1, fill_color._c
    return o


# --------------------------------------------------------------------------------------------
# Functions
# --------------------------------------------------------------------------------------------


def _in_image(image: ImageData):
    """-> (CImage, keep-alive object).  bytes are passed without a copy."""
    data = image.data
    if isinstance(data, bytes):
        keep = data
        ptr = C.cast(C.c_char_p(data), C.c_void_p)
    else:
        raw = bytes(data)
        keep = raw
        ptr = C.cast(C.c_char_p(raw), C.c_void_p)
    return CImage(image.width, image.height, ptr, len(keep)), keep


def _take(out: CImage) -> ImageData:
    try:
        data = C.string_at(out.data, out.len) if out.len else b""
        return ImageData(out.width, out.height, data)
    finally:
        library().slimg_b200_image_free(C.byref(out))


def pipeline(image: ImageData, options: CPipelineOptions, order: int = ORDER_CONVERT) -> ImageData:
    """The pixel stages of ``convert`` (crop -> extend -> resize), or the chained order."""
    cin, keep = _in_image(image)
    out = CImage()
    _check(library().slimg_b200_pipeline(C.byref(cin), C.byref(options), order, C.byref(out)))
    del keep
    return _take(out)


def crop(image: ImageData, mode: CropMode) -> ImageData:
    cin, keep = _in_image(image)
    out = CImage()
    _check(library().slimg_b200_crop(C.byref(cin), C.byref(mode._c), C.byref(out)))
    del keep
    return _take(out)


def extend(image: ImageData, mode: ExtendMode, fill: FillColor) -> ImageData:
    cin, keep = _in_image(image)
    out = CImage()
    _check(library().slimg_b200_extend(C.byref(cin), C.byref(mode._c), C.byref(fill._c), C.byref(out)))
    del keep
    return _take(out)


def resize(image: ImageData, mode: ResizeMode) -> ImageData:
    cin, keep = _in_image(image)
    out = CImage()
    _check(library().slimg_b200_resize(C.byref(cin), C.byref(mode._c), C.byref(out)))
    del keep
    return _take(out)


def calculate_dimensions(w, h, mode: ResizeMode):
    ow, oh = C.c_uint32(), C.c_uint32()
    _check(library().slimg_b200_calculate_dimensions(w, h, C.byref(mode._c), C.byref(ow), C.byref(oh)))
    return ow.value, oh.value


def resize_output_dims(w, h, mode: ResizeMode):
    ow, oh = C.c_uint32(), C.c_uint32()
    _check(library().slimg_b200_resize_output_dims(w, h, C.byref(mode._c), C.byref(ow), C.byref(oh)))
    return ow.value, oh.value


def calculate_crop_region(w, h, mode: CropMode):
    v = [C.c_uint32() for _ in range(4)]
    _check(library().slimg_b200_calculate_crop_region(w, h, C.byref(mode._c), *[C.byref(x) for x in v]))
    return tuple(x.value for x in v)


def calculate_extend_region(w, h, mode: ExtendMode):
    v = [C.c_uint32() for _ in range(4)]
    _check(library().slimg_b200_calculate_extend_region(w, h, C.byref(mode._c), *[C.byref(x) for x in v]))
    return tuple(x.value for x in v)


def pipeline_output_dims(w, h, options: CPipelineOptions, order: int = ORDER_CONVERT):
    ow, oh = C.c_uint32(), C.c_uint32()
    _check(library().slimg_b200_pipeline_output_dims(w, h, C.byref(options), order, C.byref(ow), C.byref(oh)))
    return ow.value, oh.value


def pipeline_source_window(w, h, options: CPipelineOptions, order: int = ORDER_CONVERT):
    """(x, y, width, height) of the source rectangle the pipeline reads (what host-fed calls upload)."""
    x, y, ww, hh = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
    _check(library().slimg_b200_pipeline_source_window(w, h, C.byref(options), order, C.byref(x), C.byref(y),
                                                       C.byref(ww), C.byref(hh)))
    return x.value, y.value, ww.value, hh.value


def jpeg_geometry(w, h) -> CJpegGeometry:
    g = CJpegGeometry()
    library().slimg_b200_jpeg_geometry(w, h, C.byref(g))
    return g


def jpeg_quant_table(quality: int, which: int, base: int = QT_ANNEX_K):
    out = (C.c_uint16 * 64)()
    library().slimg_b200_jpeg_quant_table(quality, which, base, out)
    return list(out)


class JpegCoefficients:
    """Coefficient planes of one image: int16 ``[HB][WB][64]`` (natural order) for Y, Cb, Cr as bytes, the
    quantisation tables they were computed with, and -- when asked for -- the un-quantised planes."""

    def __init__(self, geom, y, cb, cr, qt_luma, qt_chroma, y_uq=None, cb_uq=None, cr_uq=None):
        self.geometry, self.y, self.cb, self.cr = geom, y, cb, cr
        self.qt_luma, self.qt_chroma = list(qt_luma), list(qt_chroma)
        self.y_uq, self.cb_uq, self.cr_uq = y_uq, cb_uq, cr_uq

    @property
    def width(self): return self.geometry.width

    @property
    def height(self): return self.geometry.height


def _take_coefs(out: CJpegCoefs) -> JpegCoefficients:
    try:
        g = out.geom
        ye, ce = g.y_HB * g.y_WB * 64, g.c_HB * g.c_WB * 64
        planes = [C.string_at(out.y, ye * 2), C.string_at(out.cb, ce * 2), C.string_at(out.cr, ce * 2)]
        uq = [None, None, None]
        if out.y_uq:
            uq = [C.string_at(out.y_uq, ye * 2), C.string_at(out.cb_uq, ce * 2), C.string_at(out.cr_uq, ce * 2)]
        return JpegCoefficients(CJpegGeometry.from_buffer_copy(bytes(g)), *planes, list(out.qt_luma), list(out.qt_chroma), *uq)
    finally:
        library().slimg_b200_jpeg_coefs_free(C.byref(out))


def jpeg_frontend_ex(image: ImageData, qt_luma, qt_chroma, flags: int = 0) -> JpegCoefficients:
    cin, keep = _in_image(image)
    ql = (C.c_uint16 * 64)(*qt_luma)
    qc = (C.c_uint16 * 64)(*qt_chroma)
    out = CJpegCoefs()
    _check(library().slimg_b200_jpeg_frontend_ex(C.byref(cin), ql, qc, flags, C.byref(out)))
    del keep
    return _take_coefs(out)


def jpeg_frontend(image: ImageData, qt_luma, qt_chroma):
    """-> (geometry, y, cb, cr) with the planes as bytes of int16 [HB][WB][64]."""
    c = jpeg_frontend_ex(image, qt_luma, qt_chroma, 0)
    return c.geometry, c.y, c.cb, c.cr


def pipeline_jpeg(image: ImageData, options: CPipelineOptions, qt_luma, qt_chroma, order: int = ORDER_CONVERT,
                  flags: int = 0) -> JpegCoefficients:
    """convert(.., Jpeg) up to the entropy coder: pixel stages + JPEG front end on the device, coefficients back."""
    cin, keep = _in_image(image)
    ql = (C.c_uint16 * 64)(*qt_luma)
    qc = (C.c_uint16 * 64)(*qt_chroma)
    out = CJpegCoefs()
    _check(library().slimg_b200_pipeline_jpeg(C.byref(cin), C.byref(options), order, ql, qc, flags, C.byref(out)))
    del keep
    return _take_coefs(out)


@dataclass
class ColorPlanes:
    """Planes of one image as bytes (u8, or little-endian u16 for the 10-bit kind); ``a`` is None for an opaque WebP import."""
    width: int
    height: int
    kind: int
    chroma_width: int
    chroma_height: int
    bytes_per_sample: int
    y: bytes
    cb: bytes
    cr: bytes
    a: Optional[bytes]


def pipeline_planes(image: ImageData, options: CPipelineOptions, kind: int, order: int = ORDER_CONVERT) -> "ColorPlanes":
    """Pixel stages + the colour-plane import of the WebP / AVIF encoder, back to back on the device."""
    cin, keep = _in_image(image)
    out = CPlanes()
    _check(library().slimg_b200_pipeline_planes(C.byref(cin), C.byref(options), order, kind, C.byref(out)))
    del keep
    try:
        yb = out.width * out.height * out.bytes_per_sample
        cb = out.chroma_width * out.chroma_height * out.bytes_per_sample
        return ColorPlanes(out.width, out.height, out.kind, out.chroma_width, out.chroma_height, out.bytes_per_sample,
                           C.string_at(out.y, yb), C.string_at(out.cb, cb), C.string_at(out.cr, cb),
                           C.string_at(out.a, yb) if out.a else None)
    finally:
        library().slimg_b200_planes_free(C.byref(out))


def last_resize_path() -> int:
    return library().slimg_b200_last_resize_path()


def path_launches(path: int) -> int:
    return library().slimg_b200_path_launches(path)


def pipeline_batch_device(src_ptr: int, n: int, w: int, h: int, options: CPipelineOptions, order: int,
                          dst_ptr: int, dst_capacity: int, stream: int = 0) -> None:
    """Enqueue the pipeline for n packed device-resident images (no synchronisation)."""
    _check(library().slimg_b200_pipeline_batch_device(src_ptr, n, w, h, C.byref(options), order, dst_ptr,
                                                      dst_capacity, stream))


def jpeg_frontend_batch_device(src_ptr: int, n: int, w: int, h: int, qt_luma, qt_chroma,
                               y_ptr: int, cb_ptr: int, cr_ptr: int, stream: int = 0) -> None:
    ql = (C.c_uint16 * 64)(*qt_luma)
    qc = (C.c_uint16 * 64)(*qt_chroma)
    _check(library().slimg_b200_jpeg_frontend_batch_device(src_ptr, n, w, h, ql, qc, y_ptr, cb_ptr, cr_ptr, stream))


def jpeg_frontend_batch_device_ex(src_ptr, n, w, h, qt_luma, qt_chroma, y_ptr, cb_ptr, cr_ptr, yu_ptr, cbu_ptr, cru_ptr,
                                  stream: int = 0, flags: int = 0) -> None:
    ql = (C.c_uint16 * 64)(*qt_luma)
    qc = (C.c_uint16 * 64)(*qt_chroma)
    _check(library().slimg_b200_jpeg_frontend_batch_device_ex(src_ptr, n, w, h, ql, qc, y_ptr, cb_ptr, cr_ptr,
                                                              yu_ptr, cbu_ptr, cru_ptr, flags, stream))


def pipeline_jpeg_batch_device(src_ptr, n, w, h, options: CPipelineOptions, order, qt_luma, qt_chroma, pixels_ptr,
                               pixels_capacity, y_ptr, cb_ptr, cr_ptr, yu_ptr=None, cbu_ptr=None, cru_ptr=None,
                               stream: int = 0, flags: int = 0) -> None:
    ql = (C.c_uint16 * 64)(*qt_luma)
    qc = (C.c_uint16 * 64)(*qt_chroma)
    _check(library().slimg_b200_pipeline_jpeg_batch_device(src_ptr, n, w, h, C.byref(options), order, ql, qc, pixels_ptr,
                                                           pixels_capacity, y_ptr, cb_ptr, cr_ptr, yu_ptr, cbu_ptr, cru_ptr,
                                                           flags, stream))


def device_count() -> int:
    return library().slimg_b200_device_count()


def kernel_launches() -> int:
    return library().slimg_b200_kernel_launches()


# -- the parts of the reference's _lowlevel surface that are codec / file work (out of scope:
#    entropy coding stays on the reference libraries) ------------------------

_EXTENSIONS = {Format.JPEG: "jpg", Format.PNG: "png", Format.WEB_P: "webp", Format.AVIF: "avif",
               Format.JXL: "jxl", Format.QOI: "qoi"}


def format_extension(fmt: Format) -> str:
    return _EXTENSIONS[fmt]


def format_can_encode(fmt: Format) -> bool:
    return True   # crates/slimg-core/src/format.rs:73-75


def format_from_extension(path: str) -> Optional[Format]:
    ext = os.path.splitext(path)[1].lower().lstrip(".")
    return {"jpg": Format.JPEG, "jpeg": Format.JPEG, "png": Format.PNG, "webp": Format.WEB_P,
            "avif": Format.AVIF, "jxl": Format.JXL, "qoi": Format.QOI}.get(ext)


def format_from_magic_bytes(data: bytes) -> Optional[Format]:
    d = bytes(data[:16])
    if d[:3] == b"\xff\xd8\xff": return Format.JPEG
    if d[:8] == b"\x89PNG\r\n\x1a\n": return Format.PNG
    if d[:4] == b"RIFF" and d[8:12] == b"WEBP": return Format.WEB_P
    if d[4:12] in (b"ftypavif", b"ftypavis"): return Format.AVIF
    if d[:2] == b"\xff\x0a" or d[:12] == b"\x00\x00\x00\x0cJXL \r\n\x87\n": return Format.JXL
    if d[:4] == b"qoif": return Format.QOI
    return None


# -- convert / decode / optimize: the UniFFI-generated module exports them (crates/slimg-ffi/src/lib.rs:277-318,
#    342-351).  The pixel stages and the JPEG front end run on the GPU; entropy coding and decoding are host codecs
#    that plug in here (slimg_b200/_host_codecs.py installs what this machine has). ----------------------------------


@dataclass
class DecodeResult:
    image: ImageData
    format: Format


_ENCODERS = {}       # Format -> callable(ImageData, quality) -> bytes          (pixels in)
_JPEG_ENTROPY = []   # [callable(JpegCoefficients, quality) -> bytes]            (coefficients in: no pixel download)
_DECODERS = []       # [callable(bytes) -> Optional[DecodeResult]]
JPEG_TABLES = QT_MOZJPEG   # which base tables convert(.., JPEG) quantises with (the reference encoder is MozJPEG)


def register_encoder(fmt: Format, fn) -> None:
    _ENCODERS[fmt] = fn


def register_jpeg_entropy_coder(fn) -> None:
    _JPEG_ENTROPY[:] = [fn]


def register_decoder(fn) -> None:
    _DECODERS.append(fn)


def _load_host_codecs() -> None:
    from . import _host_codecs
    _host_codecs.install()


def convert(image: ImageData, options: PipelineOptions) -> PipelineResult:
    """slimg_core::convert (pipeline.rs:66-103): crop -> extend -> resize (one fused launch), then encode."""
    _load_host_codecs()
    fmt = options.format
    q = int(options.quality)
    copts = options._to_c()
    if fmt == Format.JPEG and _JPEG_ENTROPY:
        ql, qc = jpeg_quant_table(q, 0, JPEG_TABLES), jpeg_quant_table(q, 1, JPEG_TABLES)
        coefs = pipeline_jpeg(image, copts, ql, qc, ORDER_CONVERT)
        return PipelineResult(_JPEG_ENTROPY[0](coefs, q), fmt, coefs.width, coefs.height)
    enc = _ENCODERS.get(fmt)
    if enc is None:     # the reference reports this before touching pixels (pipeline.rs:67-69)
        raise SlimgError.Encode(f"no host encoder for {fmt.name}: entropy coding stays on the reference libraries "
                                "(slimg_b200._lowlevel.register_encoder)")
    px = pipeline(image, copts, ORDER_CONVERT)
    return PipelineResult(enc(px, q), fmt, px.width, px.height)


def decode(data: bytes) -> DecodeResult:
    _load_host_codecs()
    if format_from_magic_bytes(data) is None:
        raise SlimgError.UnknownFormat("unrecognised magic bytes")
    for dec in _DECODERS:
        res = dec(bytes(data))
        if res is not None:
            return res
    raise SlimgError.Decode("no host decoder for this format: decoding stays on the reference codecs")


def decode_file(path: str) -> DecodeResult:
    try:
        with open(path, "rb") as fh:
            data = fh.read()
    except OSError as exc:
        raise SlimgError.Io(str(exc)) from exc
    return decode(data)


def optimize(data: bytes, quality: int) -> PipelineResult:
    res = decode(data)
    return convert(res.image, PipelineOptions(format=res.format, quality=quality))


def install_as_slimg_lowlevel() -> None:
    """Make ``import slimg`` (the reference's UNMODIFIED Python package, bindings/python/slimg) run on this library:
    registers this module as ``slimg._lowlevel``, the name the wrapper imports (_types.py:13)."""
    import sys
    sys.modules["slimg._lowlevel"] = sys.modules[__name__]
