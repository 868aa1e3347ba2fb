"""ctypes loader for the CPU oracle (oracle/slimg_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package; nothing under slimg_b200/ does.  See slimg_oracle.h for what pins it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libslimg_oracle.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libljt_ref.so")

OK, ERR_ENCODE, ERR_RESIZE, ERR_CROP, ERR_EXTEND = 0, 5, 6, 7, 8


class ResizeMode(C.Structure):
    _fields_ = [("tag", C.c_int32), ("a", C.c_uint32), ("b", C.c_uint32), ("factor", C.c_double)]

    @classmethod
    def width(cls, v): return cls(1, v, 0, 0.0)
    @classmethod
    def height(cls, v): return cls(2, v, 0, 0.0)
    @classmethod
    def exact(cls, w, h): return cls(3, w, h, 0.0)
    @classmethod
    def fit(cls, w, h): return cls(4, w, h, 0.0)
    @classmethod
    def scale(cls, f): return cls(5, 0, 0, float(f))


class CropMode(C.Structure):
    _fields_ = [("tag", C.c_int32), ("x", C.c_uint32), ("y", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]

    @classmethod
    def region(cls, x, y, w, h): return cls(1, x, y, w, h)
    @classmethod
    def aspect_ratio(cls, w, h): return cls(2, 0, 0, w, h)


class ExtendMode(C.Structure):
    _fields_ = [("tag", C.c_int32), ("w", C.c_uint32), ("h", C.c_uint32)]

    @classmethod
    def aspect_ratio(cls, w, h): return cls(1, w, h)
    @classmethod
    def size(cls, w, h): return cls(2, w, h)


class JpegGeom(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("width", "height", "y_wb", "y_hb", "y_WB", "y_HB", "c_wb", "c_hb", "c_WB", "c_HB")]


class OracleError(Exception):
    def __init__(self, status: int):
        super().__init__(f"oracle status {status}")
        self.status = status


def build(force: bool = False) -> None:
    """Compile the C restatement (and, where /root/reference exists, the real libjpeg-turbo)."""
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "slimg_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    if os.path.isdir("/root/reference") and (force or not os.path.exists(_REF_PATH)):
        subprocess.check_call(["make", "-C", os.path.join(_HERE, "ref_ljt"), "-s", "-j8"])
    if os.path.isdir("/root/reference"):
        from . import stage_ref_binding
        stage_ref_binding.stage()      # the reference's own binding tests, unmodified, for tests/test_reference_binding.py


_lib: Optional[C.CDLL] = None
_ref: Optional[C.CDLL] = None
_u8p = C.POINTER(C.c_uint8)
_u16p = C.POINTER(C.c_uint16)
_i16p = C.POINTER(C.c_int16)
_u32p = C.POINTER(C.c_uint32)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.so_resize_exact.restype = C.c_int
        _lib.so_lanczos3_taps.restype = C.c_int
    return _lib


def ref_lib() -> Optional[C.CDLL]:
    """The real vendored libjpeg-turbo, or None when oracle/_ref was never built."""
    global _ref
    if _ref is None and os.path.exists(_REF_PATH):
        _ref = C.CDLL(_REF_PATH)
    return _ref


def _ptr(a: np.ndarray, t):
    return a.ctypes.data_as(t)


def _as_rgba(img: np.ndarray) -> np.ndarray:
    img = np.ascontiguousarray(img, dtype=np.uint8)
    assert img.ndim == 3 and img.shape[2] == 4, img.shape
    return img


# ---- geometry -------------------------------------------------------------------------------

def calculate_dimensions(w: int, h: int, mode: ResizeMode) -> Tuple[int, int]:
    ow, oh = C.c_uint32(), C.c_uint32()
    st = lib().so_calculate_dimensions(C.c_uint32(w), C.c_uint32(h), C.byref(mode), C.byref(ow), C.byref(oh))
    if st: raise OracleError(st)
    return ow.value, oh.value


def image_resize_dimensions(w, h, nw, nh, fill=False) -> Tuple[int, int]:
    ow, oh = C.c_uint32(), C.c_uint32()
    lib().so_image_resize_dimensions(C.c_uint32(w), C.c_uint32(h), C.c_uint32(nw), C.c_uint32(nh),
                                     C.c_int(int(fill)), C.byref(ow), C.byref(oh))
    return ow.value, oh.value


def resize_output_dims(w: int, h: int, mode: ResizeMode) -> Tuple[int, int]:
    ow, oh = C.c_uint32(), C.c_uint32()
    st = lib().so_resize_output_dims(C.c_uint32(w), C.c_uint32(h), C.byref(mode), C.byref(ow), C.byref(oh))
    if st: raise OracleError(st)
    return ow.value, oh.value


def calculate_crop_region(w: int, h: int, mode: CropMode) -> Tuple[int, int, int, int]:
    v = [C.c_uint32() for _ in range(4)]
    st = lib().so_calculate_crop_region(C.c_uint32(w), C.c_uint32(h), C.byref(mode), *[C.byref(x) for x in v])
    if st: raise OracleError(st)
    return tuple(x.value for x in v)


def calculate_extend_region(w: int, h: int, mode: ExtendMode) -> Tuple[int, int, int, int]:
    v = [C.c_uint32() for _ in range(4)]
    st = lib().so_calculate_extend_region(C.c_uint32(w), C.c_uint32(h), C.byref(mode), *[C.byref(x) for x in v])
    if st: raise OracleError(st)
    return tuple(x.value for x in v)


# ---- ops on (H, W, 4) uint8 arrays ----------------------------------------------------------

def crop(img: np.ndarray, mode: CropMode) -> np.ndarray:
    img = _as_rgba(img)
    h, w = img.shape[:2]
    x, y, cw, ch = calculate_crop_region(w, h, mode)
    out = np.empty((ch, cw, 4), np.uint8)
    lib().so_crop_pixels(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), C.c_uint32(x), C.c_uint32(y),
                         C.c_uint32(cw), C.c_uint32(ch), _ptr(out, _u8p))
    return out


def extend(img: np.ndarray, mode: ExtendMode, fill=(0, 0, 0, 0)) -> np.ndarray:
    """fill = RGBA quadruplet; FillColor::Transparent is (0,0,0,0) (extend.rs:14-21)."""
    img = _as_rgba(img)
    h, w = img.shape[:2]
    cw, ch, ox, oy = calculate_extend_region(w, h, mode)
    out = np.empty((ch, cw, 4), np.uint8)
    f = (C.c_uint8 * 4)(*fill)
    lib().so_extend_pixels(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), C.c_uint32(cw), C.c_uint32(ch),
                           C.c_uint32(ox), C.c_uint32(oy), f, _ptr(out, _u8p))
    return out


def resize_exact(img: np.ndarray, nw: int, nh: int) -> np.ndarray:
    img = _as_rgba(img)
    h, w = img.shape[:2]
    out = np.empty((nh, nw, 4), np.uint8)
    rc = lib().so_resize_exact(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), C.c_uint32(nw), C.c_uint32(nh),
                               _ptr(out, _u8p))
    if rc: raise MemoryError("oracle resize")
    return out


def resize(img: np.ndarray, mode: ResizeMode) -> np.ndarray:
    img = _as_rgba(img)
    h, w = img.shape[:2]
    nw, nh = resize_output_dims(w, h, mode)
    return resize_exact(img, nw, nh)


def pipeline(img: np.ndarray, *, crop_mode=None, extend_mode=None, fill=None, resize_mode=None) -> np.ndarray:
    """`convert` minus the encoder: crop -> extend -> resize (pipeline.rs:66-89).
    fill None with extend set = opaque white (pipeline.rs:78-80)."""
    out = _as_rgba(img)
    if crop_mode is not None:
        out = crop(out, crop_mode)
    if extend_mode is not None:
        out = extend(out, extend_mode, (255, 255, 255, 255) if fill is None else fill)
    if resize_mode is not None:
        out = resize(out, resize_mode)
    return out


def lanczos3_taps(in_len: int, out_len: int):
    m = lib().so_lanczos3_taps(C.c_uint32(in_len), C.c_uint32(out_len), None, None, None, C.c_int(0))
    left = np.empty(out_len, np.int32)
    count = np.empty(out_len, np.int32)
    w = np.empty((out_len, m), np.float32)
    lib().so_lanczos3_taps(C.c_uint32(in_len), C.c_uint32(out_len), _ptr(left, C.POINTER(C.c_int32)),
                           _ptr(count, C.POINTER(C.c_int32)), _ptr(w, C.POINTER(C.c_float)), C.c_int(m))
    return left, count, w


# ---- JPEG front end -------------------------------------------------------------------------

def jpeg_geometry(w: int, h: int) -> JpegGeom:
    g = JpegGeom()
    lib().so_jpeg_geometry(C.c_uint32(w), C.c_uint32(h), C.byref(g))
    return g


def jpeg_quant_table(quality: int, which: int, base: int = 0) -> np.ndarray:
    out = np.empty(64, np.uint16)
    lib().so_jpeg_quant_table(C.c_int(quality), C.c_int(which), C.c_int(base), _ptr(out, _u16p))
    return out


def _coef_planes(g: JpegGeom):
    return (np.zeros((g.y_HB, g.y_WB, 64), np.int16), np.zeros((g.c_HB, g.c_WB, 64), np.int16),
            np.zeros((g.c_HB, g.c_WB, 64), np.int16))


def jpeg_frontend(img: np.ndarray, qt_luma: np.ndarray, qt_chroma: np.ndarray):
    img = _as_rgba(img)
    h, w = img.shape[:2]
    y, cb, cr = _coef_planes(jpeg_geometry(w, h))
    ql = np.ascontiguousarray(qt_luma, np.uint16); qc = np.ascontiguousarray(qt_chroma, np.uint16)
    lib().so_jpeg_frontend(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), _ptr(ql, _u16p), _ptr(qc, _u16p),
                           _ptr(y, _i16p), _ptr(cb, _i16p), _ptr(cr, _i16p))
    return y, cb, cr


def jpeg_frontend_mozjpeg(img: np.ndarray, qt_luma: np.ndarray, qt_chroma: np.ndarray):
    """The planes with MozJPEG's front-end defaults: overshoot deringing, 10-bit clamp (restated, unpinned)."""
    img = _as_rgba(img)
    h, w = img.shape[:2]
    y, cb, cr = _coef_planes(jpeg_geometry(w, h))
    ql = np.ascontiguousarray(qt_luma, np.uint16); qc = np.ascontiguousarray(qt_chroma, np.uint16)
    lib().so_jpeg_frontend_mozjpeg(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), _ptr(ql, _u16p), _ptr(qc, _u16p),
                                   _ptr(y, _i16p), _ptr(cb, _i16p), _ptr(cr, _i16p))
    return y, cb, cr


def preprocess_deringing(block: np.ndarray, dc_quant: int) -> np.ndarray:
    ws = np.ascontiguousarray(block, np.int32).reshape(64).copy()
    lib().so_preprocess_deringing(ws.ctypes.data_as(C.POINTER(C.c_int32)), C.c_uint16(dc_quant))
    return ws.reshape(8, 8)


def jpeg_frontend_unquantised(img: np.ndarray):
    """The DCT output before quantisation (jfdctint.c:143-286, scaled by 8), same plane layout."""
    img = _as_rgba(img)
    h, w = img.shape[:2]
    y, cb, cr = _coef_planes(jpeg_geometry(w, h))
    lib().so_jpeg_frontend_unquantised(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), _ptr(y, _i16p), _ptr(cb, _i16p),
                                       _ptr(cr, _i16p))
    return y, cb, cr


def ref_jpeg_frontend(img: np.ndarray, qt_luma: np.ndarray, qt_chroma: np.ndarray):
    """Same planes from the REAL libjpeg-turbo (oracle/_ref); raises if it was never built."""
    r = ref_lib()
    if r is None:
        raise FileNotFoundError(_REF_PATH)
    img = _as_rgba(img)
    h, w = img.shape[:2]
    y, cb, cr = _coef_planes(jpeg_geometry(w, h))
    ql = np.ascontiguousarray(qt_luma, np.uint16); qc = np.ascontiguousarray(qt_chroma, np.uint16)
    r.ljt_frontend(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), _ptr(ql, _u16p), _ptr(qc, _u16p),
                   _ptr(y, _i16p), _ptr(cb, _i16p), _ptr(cr, _i16p))
    return y, cb, cr


def ref_jpeg_encode(img: np.ndarray, qt_luma: np.ndarray, qt_chroma: np.ndarray) -> bytes:
    """The JPEG byte stream the real libjpeg-turbo writes for this image (baseline, optimised Huffman tables)."""
    r = ref_lib()
    if r is None:
        raise FileNotFoundError(_REF_PATH)
    img = _as_rgba(img)
    h, w = img.shape[:2]
    ql = np.ascontiguousarray(qt_luma, np.uint16); qc = np.ascontiguousarray(qt_chroma, np.uint16)
    out, size = C.POINTER(C.c_ubyte)(), C.c_ulong()
    r.ljt_encode(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), _ptr(ql, _u16p), _ptr(qc, _u16p), C.byref(out), C.byref(size))
    data = C.string_at(out, size.value)
    r.ljt_free(out)
    return data


def ref_jpeg_write_coefficients(y, cb, cr, w: int, h: int, qt_luma, qt_chroma, progressive: bool = False) -> bytes:
    """hand quantised coefficients (from the CUDA front end) to the reference family's entropy
    coder with jpeg_write_coefficients (jctrans.c:43); returns the JPEG byte stream."""
    r = ref_lib()
    if r is None:
        raise FileNotFoundError(_REF_PATH)
    y = np.ascontiguousarray(y, np.int16); cb = np.ascontiguousarray(cb, np.int16); cr = np.ascontiguousarray(cr, np.int16)
    ql = np.ascontiguousarray(qt_luma, np.uint16); qc = np.ascontiguousarray(qt_chroma, np.uint16)
    out, size = C.POINTER(C.c_ubyte)(), C.c_ulong()
    r.ljt_write_coefficients(_ptr(y, _i16p), _ptr(cb, _i16p), _ptr(cr, _i16p), C.c_uint32(w), C.c_uint32(h),
                             _ptr(ql, _u16p), _ptr(qc, _u16p), C.c_int(1 if progressive else 0), C.byref(out), C.byref(size))
    data = C.string_at(out, size.value)
    r.ljt_free(out)
    return data


def ref_jpeg_decode(jpg: bytes, w: int, h: int) -> np.ndarray:
    r = ref_lib()
    if r is None:
        raise FileNotFoundError(_REF_PATH)
    out = np.empty((h, w, 4), np.uint8)
    ow, oh = C.c_uint32(), C.c_uint32()
    rc = r.ljt_decode_rgba(jpg, C.c_ulong(len(jpg)), C.byref(ow), C.byref(oh), _ptr(out, _u8p), C.c_size_t(out.nbytes))
    if rc or (ow.value, oh.value) != (w, h):
        raise ValueError(f"decoded {ow.value}x{oh.value}, expected {w}x{h}")
    return out


def ref_jpeg_quant_table(quality: int, which: int) -> np.ndarray:
    r = ref_lib()
    if r is None:
        raise FileNotFoundError(_REF_PATH)
    out = np.empty(64, np.uint16)
    r.ljt_quant_table(C.c_int(quality), C.c_int(which), _ptr(out, _u16p))
    return out


def rgb_ycc(r: int, g: int, b: int):
    o = (C.c_uint8 * 3)()
    lib().so_rgb_ycc(C.c_uint8(r), C.c_uint8(g), C.c_uint8(b), o)
    return tuple(o)


def fdct_islow(block: np.ndarray) -> np.ndarray:
    b = np.ascontiguousarray(block, np.int32).reshape(64).copy()
    lib().so_fdct_islow(_ptr(b, C.POINTER(C.c_int32)))
    return b.reshape(8, 8)


def quantize(ws: np.ndarray, qt: np.ndarray) -> np.ndarray:
    w = np.ascontiguousarray(ws, np.int32).reshape(64)
    q = np.ascontiguousarray(qt, np.uint16).reshape(64)
    out = np.empty(64, np.int16)
    lib().so_quantize(_ptr(w, C.POINTER(C.c_int32)), _ptr(q, _u16p), _ptr(out, _i16p))
    return out.reshape(8, 8)


# ---- colour-plane pre-stages of the WebP / AVIF encoders (parity unpinned: libwebp / ravif are absent) ----

def webp_yuva(img: np.ndarray):
    """libwebp's RGBA -> YUV 4:2:0 (+ A) import.  -> (y [h,w], u, v [ceil(h/2), ceil(w/2)], a [h,w] or None)."""
    img = _as_rgba(img)
    h, w = img.shape[:2]
    y = np.zeros((h, w), np.uint8); a = np.zeros((h, w), np.uint8)
    u = np.zeros(((h + 1) // 2, (w + 1) // 2), np.uint8); v = np.zeros_like(u)
    lib().so_webp_yuva.restype = C.c_int
    has_alpha = lib().so_webp_yuva(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), _ptr(y, _u8p), _ptr(u, _u8p), _ptr(v, _u8p), _ptr(a, _u8p))
    return y, u, v, (a if has_alpha else None)


def avif_planes(img: np.ndarray, depth: int = 10):
    """ravif's RGBA -> YCbCr 4:4:4 (BT.601, full range) + alpha plane, uint16 [h,w] each."""
    img = _as_rgba(img)
    h, w = img.shape[:2]
    out = [np.zeros((h, w), np.uint16) for _ in range(4)]
    lib().so_avif_planes(_ptr(img, _u8p), C.c_uint32(w), C.c_uint32(h), C.c_int(depth), *[_ptr(o, _u16p) for o in out])
    return tuple(out)
