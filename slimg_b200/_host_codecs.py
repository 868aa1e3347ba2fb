"""Host codecs behind ``_lowlevel.convert / decode``: NOT part of the GPU path.

SURVEY.md section 8 keeps entropy coding and decoding on the reference's libraries.  What this machine has:

* JPEG: ``libslimg_jpeg_handoff.so`` (slimg_b200/jpeg_handoff/) -- the libjpeg-turbo the reference vendors, driven
  through ``jpeg_write_coefficients``: it receives the coefficient planes of the CUDA front end, so a JPEG convert
  never brings pixels back to the host.  Settings as ``codec/jpeg.rs:58-64``: progressive, optimised Huffman tables.
* PNG / WebP encode and every decode: Pillow, as a stand-in for OxiPNG / libwebp / the ``image`` crate decoders,
  which are Rust crates that cannot be built here.  AVIF / JXL / QOI have no stand-in: ``SlimgError.Encode``.
"""
from __future__ import annotations

import ctypes as C
import io
import os

from . import _lowlevel as ll

_HANDOFF_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libslimg_jpeg_handoff.so")
HANDOFF_PROGRESSIVE, HANDOFF_OPTIMIZE = 1, 2
_handoff = None
_installed = False


def handoff_library():
    """libslimg_jpeg_handoff.so, or None when it was not built (``make -C slimg_b200/jpeg_handoff``)."""
    global _handoff
    if _handoff is None and os.path.exists(_HANDOFF_PATH):
        lib = C.CDLL(_HANDOFF_PATH)
        lib.slimg_jpeg_handoff_write.restype = C.c_int32
        lib.slimg_jpeg_handoff_write.argtypes = [C.POINTER(ll.CJpegCoefs), C.c_uint32, C.POINTER(C.c_void_p),
                                                 C.POINTER(C.c_size_t), C.POINTER(C.c_char_p)]
        lib.slimg_jpeg_handoff_decode.restype = C.c_int32
        lib.slimg_jpeg_handoff_decode.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.POINTER(ll.CImage),
                                                  C.POINTER(C.c_int32), C.POINTER(C.c_char_p)]
        lib.slimg_jpeg_handoff_free.argtypes = [C.c_void_p]
        _handoff = lib
    return _handoff


def jpeg_from_coefficients(coefs: "ll.JpegCoefficients", flags: int = HANDOFF_PROGRESSIVE | HANDOFF_OPTIMIZE) -> bytes:
    lib = handoff_library()
    if lib is None:
        raise ll.SlimgError.Encode("libslimg_jpeg_handoff.so is not built")
    c = ll.CJpegCoefs()
    c.geom = coefs.geometry
    c.qt_luma = (C.c_uint16 * 64)(*coefs.qt_luma)
    c.qt_chroma = (C.c_uint16 * 64)(*coefs.qt_chroma)
    keep = [C.create_string_buffer(p, len(p)) for p in (coefs.y, coefs.cb, coefs.cr)]
    c.y, c.cb, c.cr = (C.cast(k, C.POINTER(C.c_int16)) for k in keep)
    out, n, msg = C.c_void_p(), C.c_size_t(), C.c_char_p()
    rc = lib.slimg_jpeg_handoff_write(C.byref(c), flags, C.byref(out), C.byref(n), C.byref(msg))
    if rc != 0:
        raise ll.SlimgError.Encode((msg.value or b"JPEG hand-off failed").decode())
    try:
        return C.string_at(out, n.value)
    finally:
        lib.slimg_jpeg_handoff_free(out)


def _jpeg_decode(data: bytes):
    lib = handoff_library()
    if lib is None or data[:2] != b"\xff\xd8":
        return None
    img, fmt, msg = ll.CImage(), C.c_int32(), C.c_char_p()
    rc = lib.slimg_jpeg_handoff_decode(None, data, len(data), C.byref(img), C.byref(fmt), C.byref(msg))
    if rc != 0:
        raise ll.SlimgError.Decode((msg.value or b"JPEG decode failed").decode())
    try:
        return ll.DecodeResult(ll.ImageData(img.width, img.height, C.string_at(img.data, img.len)), ll.Format.JPEG)
    finally:
        lib.slimg_jpeg_handoff_free(img.data)


def _pillow_encoder(pil_name: str):
    def encode(img: "ll.ImageData", quality: int) -> bytes:
        from PIL import Image as PILImage
        pil = PILImage.frombytes("RGBA", (img.width, img.height), img.data)
        buf = io.BytesIO()
        pil.save(buf, pil_name, quality=max(1, quality))
        return buf.getvalue()
    return encode


def _pillow_jpeg(img: "ll.ImageData", quality: int) -> bytes:
    from PIL import Image as PILImage
    pil = PILImage.frombytes("RGBA", (img.width, img.height), img.data).convert("RGB")   # alpha dropped, codec/mod.rs:42-48
    buf = io.BytesIO()
    pil.save(buf, "JPEG", quality=max(1, quality))
    return buf.getvalue()


def _pillow_decode(data: bytes):
    fmt = ll.format_from_magic_bytes(data)
    try:
        from PIL import Image as PILImage
        pil = PILImage.open(io.BytesIO(data)).convert("RGBA")
    except ImportError:
        return None
    except Exception as exc:
        raise ll.SlimgError.Decode(str(exc)) from exc
    return ll.DecodeResult(ll.ImageData(pil.width, pil.height, pil.tobytes()), fmt)


def install() -> None:
    global _installed
    if _installed:
        return
    _installed = True
    if handoff_library() is not None:
        ll.register_jpeg_entropy_coder(lambda coefs, q: jpeg_from_coefficients(coefs))
        ll.register_decoder(_jpeg_decode)
    try:
        import PIL  # noqa: F401
    except ImportError:
        return
    ll._ENCODERS.setdefault(ll.Format.PNG, _pillow_encoder("PNG"))
    ll._ENCODERS.setdefault(ll.Format.WEB_P, _pillow_encoder("WEBP"))
    ll._ENCODERS.setdefault(ll.Format.JPEG, _pillow_jpeg)     # only used when the hand-off library is absent
    ll.register_decoder(_pillow_decode)
