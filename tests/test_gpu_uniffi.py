"""The pixel functions through the UniFFI scaffolding symbols (what the unmodified Kotlin / Python bindings call),
on the GPU, against the oracle.  Cases follow bindings/python/tests/test_operations.py and SlimgTest.kt."""
import ctypes as C
import struct

import numpy as np
import pytest

import oracle
import uniffi_client as U
from conftest import binding_image, noise_image

pytestmark = pytest.mark.gpu


def _img(a):
    return U.w_image(a.shape[1], a.shape[0], a.tobytes())


def _lift(b):
    (w, h, data), off = U.r_image(b)
    assert off == len(b)
    return np.frombuffer(data, np.uint8).reshape(h, w, 4)


def test_crop_and_extend_bit_exact_through_uniffi():
    a = binding_image(100, 80)
    got = _lift(U.call("crop", _img(a), U.w_crop("region", 10, 20, 30, 40)))           # test_operations.py:12-16
    assert np.array_equal(got, oracle.crop(a, oracle.CropMode.region(10, 20, 30, 40)))
    got = _lift(U.call("crop", _img(a), U.w_crop("aspect", 16, 9)))
    assert np.array_equal(got, oracle.crop(a, oracle.CropMode.aspect_ratio(16, 9)))
    got = _lift(U.call("extend", _img(a), U.w_extend("aspect", 1, 1), U.w_fill((255, 0, 0, 255))))   # SlimgTest.kt:181-227
    assert np.array_equal(got, oracle.extend(a, oracle.ExtendMode.aspect_ratio(1, 1), (255, 0, 0, 255)))
    got = _lift(U.call("extend", _img(a), U.w_extend("size", 120, 90), U.w_fill()))
    assert np.array_equal(got, oracle.extend(a, oracle.ExtendMode.size(120, 90), (0, 0, 0, 0)))


@pytest.mark.parametrize("mode,lowered", [
    (oracle.ResizeMode.width(50), U.w_resize("width", 50)),
    (oracle.ResizeMode.height(20), U.w_resize("height", 20)),
    (oracle.ResizeMode.exact(20, 20), U.w_resize("exact", 20, 20)),
    (oracle.ResizeMode.fit(40, 40), U.w_resize("fit", 40, 40)),
    (oracle.ResizeMode.scale(2.0), U.w_resize("scale", 2.0)),
])
def test_resize_through_uniffi(mode, lowered):
    a = noise_image(100, 80, 3)
    got = _lift(U.call("resize", _img(a), lowered))
    want = oracle.resize(a, mode)
    assert got.shape == want.shape
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 1, f"max abs err {d.max()}"          # +-1 LSB per u8 channel (BASELINE.json)


def test_convert_runs_the_pipeline_then_the_encoder_hook():
    libc = C.CDLL(None)
    libc.malloc.restype, libc.malloc.argtypes = C.c_void_p, [C.c_size_t]
    seen = {}

    class CImage(C.Structure):
        _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("data", C.POINTER(C.c_uint8)), ("len", C.c_size_t)]

    def encode(user, pixels, fmt, quality, out, out_len, message):      # stands in for the reference's codec
        px = C.cast(pixels, C.POINTER(CImage)).contents
        seen.update(fmt=fmt, quality=quality, w=px.width, h=px.height)
        payload = b"RAW0" + C.string_at(px.data, px.len)
        p = libc.malloc(len(payload))
        C.memmove(p, payload, len(payload))
        out[0], out_len[0] = p, len(payload)
        return 0

    hook = U.ENCODE_HOOK(encode)
    U.lib().slimg_b200_set_codec_hooks(C.cast(hook, C.c_void_p), None, None)
    try:
        a = noise_image(120, 90, 5)
        # SlimgTest.kt:324-341 / test_convert.py:55-65: crop 16:9 then extend 1:1 -> square; default fill is opaque white
        opts = U.w_options(2, 77, resize=U.w_resize("width", 60), crop=U.w_crop("aspect", 16, 9), extend=U.w_extend("aspect", 1, 1))
        data, fmt, w, h = U.r_pipeline_result(U.call("convert", _img(a), opts))
        want = oracle.pipeline(a, crop_mode=oracle.CropMode.aspect_ratio(16, 9), extend_mode=oracle.ExtendMode.aspect_ratio(1, 1),
                               fill=(255, 255, 255, 255), resize_mode=oracle.ResizeMode.width(60))
        assert (fmt, w, h) == (2, want.shape[1], want.shape[0]) and w == h == 60      # pipeline.rs:97-102: post-resize dims
        assert seen == {"fmt": 2, "quality": 77, "w": 60, "h": 60}
        got = np.frombuffer(data[4:], np.uint8).reshape(h, w, 4)
        assert data[:4] == b"RAW0" and np.abs(got.astype(int) - want.astype(int)).max() <= 1
    finally:
        U.lib().slimg_b200_set_codec_hooks(None, None, None)
    with pytest.raises(U.UniffiError) as e:                  # no encoder linked -> SlimgError::Encode, after the pixel stages ran
        U.call("convert", _img(noise_image(16, 16)), U.w_options(1, 80))
    assert e.value.variant == "Encode"


@pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")
def test_convert_jpeg_through_the_coefficient_hook():
    """Format::Jpeg with a coefficient-taking entropy coder installed (include/slimg_uniffi.h): convert runs
    crop -> extend -> resize -> JPEG front end on the device and hands the planes to libjpeg-turbo's
    jpeg_write_coefficients (slimg_b200/jpeg_handoff) -- the file equals the full CPU encoder's, byte for byte."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(U._build.LIB_PATH)), "libslimg_jpeg_handoff.so")
    if not os.path.exists(path):
        pytest.skip("libslimg_jpeg_handoff.so not built")
    handoff = C.CDLL(path)
    hook = C.cast(handoff.slimg_jpeg_handoff_hook, C.c_void_p)
    L = U.lib()
    L.slimg_b200_set_jpeg_coefficient_hook.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint32]
    L.slimg_b200_kernel_launches.restype = C.c_uint64
    # user = flags + 1: baseline scan script, optimised Huffman tables (what oracle.ref_jpeg_encode uses); Annex-K tables
    L.slimg_b200_set_jpeg_coefficient_hook(hook, C.c_void_p(2 + 1), 0, 0)
    try:
        a = noise_image(401, 299, 71)
        ql, qc = oracle.ref_jpeg_quant_table(70, 0), oracle.ref_jpeg_quant_table(70, 1)
        opts = U.w_options(1, 70, crop=U.w_crop("region", 7, 9, 300, 200), extend=U.w_extend("size", 333, 257), fill=U.w_fill((10, 200, 30, 128)))
        before = L.slimg_b200_kernel_launches()
        data, fmt, w, h = U.r_pipeline_result(U.call("convert", _img(a), opts))
        assert L.slimg_b200_kernel_launches() - before == 2          # K1 + K3; no pixel download in between
        want_px = oracle.pipeline(a, crop_mode=oracle.CropMode.region(7, 9, 300, 200), extend_mode=oracle.ExtendMode.size(333, 257),
                                  fill=(10, 200, 30, 128))
        assert (fmt, w, h) == (1, 333, 257) and data == oracle.ref_jpeg_encode(want_px, ql, qc)
        # with a resize in front: the file is a JPEG of the post-resize size (pipeline.rs:97-102) that decodes to the oracle's
        # pixels within JPEG loss
        opts = U.w_options(1, 90, resize=U.w_resize("width", 160), crop=U.w_crop("aspect", 16, 9), extend=U.w_extend("aspect", 1, 1))
        data, fmt, w, h = U.r_pipeline_result(U.call("convert", _img(gradient := np.ascontiguousarray(np.tile(np.arange(400, dtype=np.uint8)[None, :, None] // 2, (300, 1, 4)))), opts))
        assert (fmt, w, h) == (1, 160, 160) and data[:2] == b"\xff\xd8"
        dec = oracle.ref_jpeg_decode(data, 160, 160)
        want = oracle.pipeline(gradient, crop_mode=oracle.CropMode.aspect_ratio(16, 9), extend_mode=oracle.ExtendMode.aspect_ratio(1, 1),
                               fill=(255, 255, 255, 255), resize_mode=oracle.ResizeMode.width(160))
        assert np.abs(dec[..., :3].astype(int) - want[..., :3].astype(int)).mean() < 3.0
    finally:
        L.slimg_b200_set_jpeg_coefficient_hook(None, None, 1, 0)
