"""The UniFFI scaffolding symbols of libslimg_ffi, exported by libslimg_b200.so (include/slimg_uniffi.h): symbol
set, contract version, checksums, buffer management, wire format of the host-only helpers and of the error path.
Expected values are the reference's: slimg_ffi.kt:662-706, 809-852; format.rs tests; pipeline.rs:148-162."""
import ctypes as C
import os
import re
import struct

import pytest

import uniffi_client as U
from slimg_b200 import _build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_uniffi_symbol_is_exported():
    text = open(os.path.join(ROOT, "include", "slimg_uniffi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = sorted(set(re.findall(r"\b((?:uniffi_slimg_ffi|ffi_slimg_ffi|slimg_b200)_[a-z0-9_]+)\s*\(", text)))
    assert len(names) == 4 + 1 + 12 + 12 + 2      # + the two hook setters
    lib = C.CDLL(_build.build())
    for n in names:
        assert hasattr(lib, n), n
    # the names the generated Kotlin binding registers (slimg_ffi.kt:675-706)
    for fn in ("convert", "crop", "decode", "decode_file", "extend", "format_can_encode", "format_extension",
               "format_from_extension", "format_from_magic_bytes", "optimize", "output_path", "resize"):
        assert f"uniffi_slimg_ffi_fn_func_{fn}" in names and f"uniffi_slimg_ffi_checksum_func_{fn}" in names


def test_contract_version_and_checksums():
    lib = U.lib()
    assert lib.ffi_slimg_ffi_uniffi_contract_version() == 30                      # slimg_ffi.kt:809
    want = {"convert": 14232, "crop": 14284, "decode": 62226, "decode_file": 64209, "extend": 52877,
            "format_can_encode": 31855, "format_extension": 23550, "format_from_extension": 56933,
            "format_from_magic_bytes": 41284, "optimize": 28797, "output_path": 60410, "resize": 48866}
    for fn, v in want.items():                                                    # slimg_ffi.kt:818-852
        f = getattr(lib, f"uniffi_slimg_ffi_checksum_func_{fn}")
        f.restype = C.c_uint16
        assert f() == v, fn


def test_rustbuffer_alloc_reserve_free():
    lib, st = U.lib(), U.RustCallStatus()
    b = lib.ffi_slimg_ffi_rustbuffer_alloc(100, C.byref(st))
    assert st.code == 0 and b.capacity >= 100 and b.len == 100 and bool(b.data)
    assert bytes(b.data[:100]) == bytes(100)
    b.data[0] = 7
    b.len = 1
    b2 = lib.ffi_slimg_ffi_rustbuffer_reserve(b, 4096, C.byref(st))
    assert st.code == 0 and b2.capacity >= 4097 and b2.len == 1 and b2.data[0] == 7
    lib.ffi_slimg_ffi_rustbuffer_free(b2, C.byref(st))
    assert st.code == 0
    assert U.from_buffer(U.to_buffer(b"hello")) == b"hello"
    assert U.from_buffer(U.to_buffer(b"")) == b""


def test_format_helpers_match_format_rs():
    ext = {1: "jpg", 2: "png", 3: "webp", 4: "avif", 5: "jxl", 6: "qoi"}          # format.rs:58-67
    for f, e in ext.items():
        assert U.call("format_extension", U.w_format(f)) == e.encode()
        st = U.RustCallStatus()
        assert U.lib().uniffi_slimg_ffi_fn_func_format_can_encode(U.to_buffer(U.w_format(f)), C.byref(st)) == 1
    some = lambda f: b"\x01" + struct.pack(">i", f)
    cases = {"photo.jpg": 1, "photo.jpeg": 1, "photo.JPG": 1, "photo.JpEg": 1, "image.png": 2, "image.PNG": 2,
             "image.webp": 3, "image.avif": 4, "image.jxl": 5, "image.qoi": 6, "/some/dir.d/photo.WebP": 3}
    for path, f in cases.items():                                                 # format.rs:84-170
        assert U.call("format_from_extension", path.encode()) == some(f), path
    for path in ("file.bmp", "noextension", ".jpg", "dir.jpg/file"):
        assert U.call("format_from_extension", path.encode()) == b"\x00", path
    magic = {b"\xff\xd8\xff\xe0" + bytes(8): 1, b"\x89PNG\r\n\x1a\n": 2, b"RIFF\0\0\0\0WEBP": 3,
             b"\0\0\0\x20ftypavif": 4, b"\0\0\0\x20ftypavis": 4, b"\xff\x0a": 5, b"\0\0\0\x0cJXL \r\n": 5, b"qoif": 6}
    for data, f in magic.items():                                                 # format.rs:30-55
        assert U.call("format_from_magic_bytes", U.w_bytes(data)) == some(f), data
    for data in (b"", b"\xff", b"RIFF\0\0\0\0WAVE", b"\0\0\0\x20ftypheic"):
        assert U.call("format_from_magic_bytes", U.w_bytes(data)) == b"\x00", data


def test_output_path_matches_pipeline_rs(tmp_path):
    none = U.w_option(None)
    assert U.call("output_path", b"/tmp/photo.jpg", U.w_format(3), none) == b"/tmp/photo.webp"      # pipeline.rs:148-152
    assert U.call("output_path", b"/tmp/photo.jpg", U.w_format(2), U.w_option(U.w_string("/out/result.png"))) == b"/out/result.png"
    d = str(tmp_path)
    assert U.call("output_path", b"/tmp/photo.jpg", U.w_format(4), U.w_option(U.w_string(d))) == f"{d}/photo.avif".encode()
    assert U.call("output_path", b"noext", U.w_format(6), none) == b"noext.qoi"


def test_errors_are_lowered_as_slimg_error():
    img = U.w_image(10, 8, bytes(10 * 8 * 4))
    with pytest.raises(U.UniffiError) as e:                                       # crop.rs:28-37: region out of bounds
        U.call("crop", img, U.w_crop("region", 8, 0, 5, 4))
    assert e.value.variant == "Crop" and e.value.message
    with pytest.raises(U.UniffiError) as e:                                       # extend.rs:66-73: smaller than the image
        U.call("extend", img, U.w_extend("size", 5, 8), U.w_fill())
    assert e.value.variant == "Extend"
    with pytest.raises(U.UniffiError) as e:                                       # resize.rs:46-52: scale <= 0
        U.call("resize", img, U.w_resize("scale", 0.0))
    assert e.value.variant == "Resize"
    with pytest.raises(U.UniffiError) as e:                                       # pipeline.rs:71-74: crop fails first
        U.call("convert", img, U.w_options(1, 80, resize=U.w_resize("width", 5), crop=U.w_crop("region", 0, 0, 50, 4)))
    assert e.value.variant == "Crop"
    with pytest.raises(U.UniffiError) as e:                                       # no decoder linked: Decode, like a bad file
        U.call("decode", U.w_bytes(b"\xff\xd8\xff\xe0"))
    assert e.value.variant == "Decode"
    with pytest.raises(U.UniffiError) as e:
        U.call("decode_file", b"/nonexistent/file.jpg")
    assert e.value.variant == "Io"
    with pytest.raises(U.UniffiPanic):                                            # a buffer that does not parse
        U.call("resize", img[:-3], U.w_resize("width", 5))
    with pytest.raises(U.UniffiPanic):
        U.call("crop", img, struct.pack(">i", 9))
