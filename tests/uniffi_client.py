"""A minimal UniFFI foreign-side client over ctypes: lowers arguments into RustBuffers and lifts results exactly as
the reference's generated Kotlin binding does (bindings/kotlin/.../slimg_ffi.kt: big-endian, RustBuffer by value,
RustCallStatus out-parameter).  Test infrastructure: it drives the uniffi_slimg_ffi_* symbols of libslimg_b200.so."""
import ctypes as C
import struct

from slimg_b200 import _build


class RustBuffer(C.Structure):
    _fields_ = [("capacity", C.c_uint64), ("len", C.c_uint64), ("data", C.POINTER(C.c_uint8))]


class ForeignBytes(C.Structure):
    _fields_ = [("len", C.c_int32), ("data", C.POINTER(C.c_uint8))]


class RustCallStatus(C.Structure):
    _fields_ = [("code", C.c_int8), ("error_buf", RustBuffer)]


ERROR_NAMES = {1: "UnsupportedFormat", 2: "UnknownFormat", 3: "EncodingNotSupported", 4: "Decode", 5: "Encode",
               6: "Resize", 7: "Crop", 8: "Extend", 9: "Io", 10: "Image"}


class UniffiError(Exception):
    def __init__(self, tag, message):
        super().__init__(f"{ERROR_NAMES.get(tag, tag)}: {message}")
        self.tag, self.variant, self.message = tag, ERROR_NAMES.get(tag), message


class UniffiPanic(Exception):
    pass


ENCODE_HOOK = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_uint8, C.POINTER(C.c_void_p),
                          C.POINTER(C.c_size_t), C.POINTER(C.c_char_p))
DECODE_HOOK = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.POINTER(C.c_uint8), C.c_size_t, C.c_void_p, C.POINTER(C.c_int32),
                          C.POINTER(C.c_char_p))

_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_build.build())
        st = C.POINTER(RustCallStatus)
        L.ffi_slimg_ffi_rustbuffer_alloc.restype, L.ffi_slimg_ffi_rustbuffer_alloc.argtypes = RustBuffer, [C.c_uint64, st]
        L.ffi_slimg_ffi_rustbuffer_from_bytes.restype, L.ffi_slimg_ffi_rustbuffer_from_bytes.argtypes = RustBuffer, [ForeignBytes, st]
        L.ffi_slimg_ffi_rustbuffer_free.restype, L.ffi_slimg_ffi_rustbuffer_free.argtypes = None, [RustBuffer, st]
        L.ffi_slimg_ffi_rustbuffer_reserve.restype, L.ffi_slimg_ffi_rustbuffer_reserve.argtypes = RustBuffer, [RustBuffer, C.c_uint64, st]
        L.ffi_slimg_ffi_uniffi_contract_version.restype = C.c_uint32
        for name, nargs in (("resize", 2), ("crop", 2), ("extend", 3), ("convert", 2), ("decode", 1), ("decode_file", 1),
                            ("format_extension", 1), ("format_from_extension", 1), ("format_from_magic_bytes", 1),
                            ("output_path", 3)):
            fn = getattr(L, f"uniffi_slimg_ffi_fn_func_{name}")
            fn.restype, fn.argtypes = RustBuffer, [RustBuffer] * nargs + [st]
        L.uniffi_slimg_ffi_fn_func_optimize.restype = RustBuffer
        L.uniffi_slimg_ffi_fn_func_optimize.argtypes = [RustBuffer, C.c_uint8, st]
        L.uniffi_slimg_ffi_fn_func_format_can_encode.restype = C.c_int8
        L.uniffi_slimg_ffi_fn_func_format_can_encode.argtypes = [RustBuffer, st]
        L.slimg_b200_set_codec_hooks.restype, L.slimg_b200_set_codec_hooks.argtypes = None, [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


# ---- lowering (what FfiConverter*.write does) --------------------------------------------------------------------

def to_buffer(payload: bytes) -> RustBuffer:
    st = RustCallStatus()
    arr = (C.c_uint8 * len(payload)).from_buffer_copy(payload) if payload else None
    fb = ForeignBytes(len(payload), C.cast(arr, C.POINTER(C.c_uint8)) if payload else None)
    buf = lib().ffi_slimg_ffi_rustbuffer_from_bytes(fb, C.byref(st))
    assert st.code == 0
    return buf


def from_buffer(buf: RustBuffer) -> bytes:
    data = C.string_at(buf.data, buf.len) if buf.len else b""
    st = RustCallStatus()
    lib().ffi_slimg_ffi_rustbuffer_free(buf, C.byref(st))
    return data


def w_image(width, height, data: bytes) -> bytes:
    return struct.pack(">IIi", width, height, len(data)) + data


def w_resize(kind, *v) -> bytes:
    if kind in ("width", "height"):
        return struct.pack(">iI", 1 if kind == "width" else 2, v[0])
    if kind in ("exact", "fit"):
        return struct.pack(">iII", 3 if kind == "exact" else 4, *v)
    return struct.pack(">id", 5, v[0])


def w_crop(kind, *v) -> bytes:
    return struct.pack(">iIIII", 1, *v) if kind == "region" else struct.pack(">iII", 2, *v)


def w_extend(kind, w, h) -> bytes:
    return struct.pack(">iII", 1 if kind == "aspect" else 2, w, h)


def w_fill(rgba=None) -> bytes:
    return struct.pack(">i", 2) if rgba is None else struct.pack(">iBBBB", 1, *rgba)


def w_format(f: int) -> bytes:
    return struct.pack(">i", f)


def w_option(payload) -> bytes:
    return b"\x00" if payload is None else b"\x01" + payload


def w_options(fmt, quality, resize=None, crop=None, extend=None, fill=None) -> bytes:
    return w_format(fmt) + struct.pack(">B", quality) + w_option(resize) + w_option(crop) + w_option(extend) + w_option(fill)


def w_bytes(b: bytes) -> bytes:
    return struct.pack(">i", len(b)) + b


def w_string(s: str) -> bytes:
    return w_bytes(s.encode())


# ---- lifting -----------------------------------------------------------------------------------------------------

def r_image(b: bytes, off=0):
    w, h, n = struct.unpack_from(">IIi", b, off)
    return (w, h, b[off + 12:off + 12 + n]), off + 12 + n


def r_pipeline_result(b: bytes):
    n, = struct.unpack_from(">i", b, 0)
    data = b[4:4 + n]
    fmt, w, h = struct.unpack_from(">iII", b, 4 + n)
    assert 4 + n + 12 == len(b)
    return data, fmt, w, h


def _check(st: RustCallStatus):
    if st.code == 0:
        return
    payload = from_buffer(st.error_buf)
    if st.code == 1:
        tag, n = struct.unpack_from(">ii", payload, 0)
        raise UniffiError(tag, payload[8:8 + n].decode())
    raise UniffiPanic(payload.decode(errors="replace"))


def call(name, *payloads, raw_strings=()):
    """payloads: lowered argument bytes; returns the lifted result bytes (RustBuffer content)."""
    st = RustCallStatus()
    bufs = [to_buffer(p) for p in payloads]
    ret = getattr(lib(), f"uniffi_slimg_ffi_fn_func_{name}")(*bufs, C.byref(st))
    _check(st)
    return from_buffer(ret)
