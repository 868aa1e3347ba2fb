"""Round-2 GPU parity cases (all through the C ABI): every K2 kernel path is executed and named, the JPEG front end as
a stage of convert (pixels stay in HBM), the un-quantised side output, the hand-off to the reference library's entropy
coder, mixed-size and pageable batches, the plan cache under churn, tall canvases."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle
from conftest import noise_image, gradient_image, constant_image
from slimg_b200 import _lowlevel as ll
from test_gpu_parity import _close, _ll_img, _np, _modes, _planes, WHITE, CLEAR

pytestmark = pytest.mark.gpu


# ---- every K2 path, by name ---------------------------------------------------------------------------------------

PATH_CASES = [
    # (label, w, h, resize, crop, extend, fill, order)
    ("down5", 1200, 800, ("width", 240), None, None, None, "convert"),
    ("c3-like chain", 1500, 1000, ("width", 200), ("aspect", 16, 9), ("aspect", 1, 1), WHITE, "chain"),
    ("c3-like convert", 1500, 1000, ("width", 200), ("aspect", 16, 9), ("aspect", 1, 1), WHITE, "convert"),
    ("down16 (bands beyond the V-first ring)", 4000, 3000, ("exact", 250, 188), None, None, None, "convert"),
    ("portrait, side padding takes part in the taps", 1000, 3000, ("width", 400), None, ("aspect", 16, 9), (7, 99, 201, 255), "convert"),
    ("odd crop x on a 1200-wide source", 1200, 900, ("width", 300), ("region", 7, 5, 1100, 800), None, None, "convert"),
    ("down10 (5-chunk bands: the GUI thumbnail of a 12 MP photo)", 4000, 3000, ("fit", 400, 400), None, None, None, "convert"),
    ("down11 with crop and padding", 3300, 2200, ("width", 300), ("aspect", 4, 3), ("aspect", 1, 1), (9, 8, 7, 255), "chain"),
    ("down2", 1024, 768, ("scale", 0.5), None, None, None, "convert"),
    ("down1.5", 900, 600, ("width", 600), None, None, None, "convert"),
]
ALLOWED = {
    "vfirst": {ll.PATH_VFIRST, ll.PATH_HFIRST, ll.PATH_GENERIC, ll.PATH_TILE},
    "tc": {ll.PATH_TC, ll.PATH_VFIRST, ll.PATH_HFIRST, ll.PATH_GENERIC, ll.PATH_TILE},
    "hfirst": {ll.PATH_HFIRST, ll.PATH_GENERIC, ll.PATH_TILE},
    "generic": {ll.PATH_GENERIC, ll.PATH_FEWTAPS},
}


def _want(img, o, fill, order):
    if order == "convert":
        return oracle.pipeline(img, **o)
    want = img
    if o["resize_mode"] is not None: want = oracle.resize(want, o["resize_mode"])
    if o["crop_mode"] is not None: want = oracle.crop(want, o["crop_mode"])
    if o["extend_mode"] is not None: want = oracle.extend(want, o["extend_mode"], fill if fill is not None else WHITE)
    return want


@pytest.mark.parametrize("path", ["vfirst", "tc", "hfirst", "generic"])
@pytest.mark.parametrize("label,w,h,resize,crop,extend,fill,order", PATH_CASES, ids=[c[0] for c in PATH_CASES])
def test_every_resize_path(monkeypatch, path, label, w, h, resize, crop, extend, fill, order):
    monkeypatch.setenv("SLIMG_B200_RESIZE_PATH", path)
    img = noise_image(w, h, 61)
    o, l = _modes(resize, crop, extend, fill)
    got = _np(ll.pipeline(_ll_img(img), l, ll.ORDER_CHAIN if order == "chain" else ll.ORDER_CONVERT))
    used = ll.last_resize_path()
    assert used in ALLOWED[path], (path, ll.PATH_NAMES[used])
    _close(got, _want(img, o, fill, order), "noise", f"path={path} {label}")


def test_default_paths_are_the_intended_kernels(monkeypatch):
    monkeypatch.delenv("SLIMG_B200_RESIZE_PATH", raising=False)
    img = noise_image(1200, 800, 62)
    ll.resize(_ll_img(img), ll.ResizeMode.WIDTH(240))
    assert ll.last_resize_path() in (ll.PATH_VFIRST, ll.PATH_TC)          # the integer tensor pipeline
    ll.crop(_ll_img(img), ll.CropMode.ASPECT_RATIO(1, 1))
    assert ll.last_resize_path() == ll.PATH_REMAP
    ll.resize(_ll_img(noise_image(64, 64, 1)), ll.ResizeMode.EXACT(200, 20))      # up on x, down on y: one tile holds the footprint
    assert ll.last_resize_path() == ll.PATH_TILE
    ll.resize(_ll_img(noise_image(1024, 1024, 1)), ll.ResizeMode.SCALE(4.0))       # BASELINE C4b's shape
    assert ll.last_resize_path() == ll.PATH_TILE
    ll.resize(_ll_img(noise_image(63, 64, 1)), ll.ResizeMode.EXACT(200, 20))      # rows that are not 16-byte aligned: re-pitched, then the same kernel
    assert ll.last_resize_path() == ll.PATH_TILE
    ll.resize(_ll_img(noise_image(1199, 800, 62)), ll.ResizeMode.WIDTH(240))
    assert ll.last_resize_path() in (ll.PATH_VFIRST, ll.PATH_TC)
    monkeypatch.setenv("SLIMG_B200_NO_REPITCH", "1")                               # without the re-pitch pass: the f32 kernels
    ll.resize(_ll_img(noise_image(63, 60, 1)), ll.ResizeMode.EXACT(200, 20))
    assert ll.last_resize_path() in (ll.PATH_GENERIC, ll.PATH_FEWTAPS)
    ll.resize(_ll_img(noise_image(4000, 3000, 1)), ll.ResizeMode.FIT(400, 400))   # 10x (the GUI's thumbnail of a 12 MP photo): 5-chunk bands
    assert ll.last_resize_path() == ll.PATH_VFIRST
    ll.resize(_ll_img(noise_image(6000, 4000, 1)), ll.ResizeMode.FIT(400, 400))   # 15x: bands beyond the V-first ring
    assert ll.last_resize_path() == ll.PATH_HFIRST
    before = ll.path_launches(ll.PATH_REMAP)
    ll.extend(_ll_img(img), ll.ExtendMode.ASPECT_RATIO(1, 1), ll.FillColor.TRANSPARENT())
    assert ll.path_launches(ll.PATH_REMAP) == before + 1


@pytest.mark.parametrize("kind,v", [("constant", 0), ("constant", 128), ("constant", 255)])
@pytest.mark.parametrize("path", ["vfirst", "tc", "generic"])
def test_constant_images_come_back_exactly(monkeypatch, path, kind, v):
    monkeypatch.setenv("SLIMG_B200_RESIZE_PATH", path)
    img = constant_image(800, 600, v)
    got = _np(ll.resize(_ll_img(img), ll.ResizeMode.WIDTH(160)))
    assert got.shape == (120, 160, 4) and (got == v).all()


# ---- K1 beyond gridDim.y ---------------------------------------------------------------------------------------------

def test_extend_to_a_canvas_taller_than_262140_rows():
    img = noise_image(4, 6, 63)
    got = _np(ll.extend(_ll_img(img), ll.ExtendMode.SIZE(4, 262151), ll.FillColor.SOLID(1, 2, 3, 4)))
    want = oracle.extend(img, oracle.ExtendMode.size(4, 262151), (1, 2, 3, 4))
    assert np.array_equal(got, want)


# ---- JPEG front end as a stage of convert ---------------------------------------------------------------------------------

def _coef_planes(c):
    return _planes(c.geometry, c.y, c.cb, c.cr)


@pytest.mark.parametrize("w,h,resize,crop,extend,fill", [
    (600, 400, ("width", 250), ("aspect", 16, 9), ("aspect", 1, 1), WHITE),
    (401, 299, None, ("region", 7, 9, 300, 200), ("size", 333, 257), (10, 200, 30, 128)),
    (1203, 807, ("fit", 400, 400), None, None, None),
    (640, 480, None, None, None, None),
])
def test_pipeline_jpeg_is_the_front_end_of_the_pipelines_pixels(w, h, resize, crop, extend, fill):
    """convert(.., Jpeg) on the device: coefficients of pipeline_jpeg == front end of the pixels `pipeline` returns,
    == the oracle's front end of those pixels (pipeline.rs:86-95 -> codec/jpeg.rs:51-72)."""
    img = noise_image(w, h, 64)
    o, l = _modes(resize, crop, extend, fill)
    ql, qc = ll.jpeg_quant_table(70, 0), ll.jpeg_quant_table(70, 1)
    before = ll.kernel_launches()
    coefs = ll.pipeline_jpeg(_ll_img(img), l, ql, qc, ll.ORDER_CONVERT)
    launched = ll.kernel_launches() - before
    repitch = 1 if resize is not None and w % 4 else 0           # rows that are not 16-byte aligned: one K1 copy into a padded pitch first
    assert launched == (1 if (resize, crop, extend) == (None, None, None) else 2 + repitch)        # pixel stages + K3, nothing else
    px = ll.pipeline(_ll_img(img), l, ll.ORDER_CONVERT)
    assert (coefs.width, coefs.height) == (px.width, px.height)
    want = oracle.jpeg_frontend(_np(px), np.array(ql, np.uint16), np.array(qc, np.uint16))
    for g, wv, name in zip(_coef_planes(coefs), want, ("y", "cb", "cr")):
        assert np.array_equal(g, wv), f"{name}: {np.count_nonzero(g != wv)} coefficients differ"
    if resize is None:            # crop / extend are bit-exact, so the whole convert is pinned to the oracle's pixels too
        want2 = oracle.jpeg_frontend(oracle.pipeline(img, **o) if (crop or extend) else img, np.array(ql, np.uint16), np.array(qc, np.uint16))
        for g, wv in zip(_coef_planes(coefs), want2):
            assert np.array_equal(g, wv)


@pytest.mark.parametrize("w,h", [(16, 16), (53, 37), (641, 480), (1920, 1080)])
def test_unquantised_side_output_is_the_islow_dct(w, h):
    """The optional second output (MozJPEG's trellis input): bit-equal to jpeg_fdct_islow (jfdctint.c:143-286) of the
    level-shifted samples; the quantised planes are unchanged by asking for it."""
    img = noise_image(w, h, 65)
    ql, qc = ll.jpeg_quant_table(70, 0, ll.QT_MOZJPEG), ll.jpeg_quant_table(70, 1, ll.QT_MOZJPEG)
    c = ll.jpeg_frontend_ex(_ll_img(img), ql, qc, ll.JPEG_UNQUANTISED)
    plain = ll.jpeg_frontend_ex(_ll_img(img), ql, qc, 0)
    assert plain.y_uq is None and (c.y, c.cb, c.cr) == (plain.y, plain.cb, plain.cr)
    want = oracle.jpeg_frontend_unquantised(img)
    got = _planes(c.geometry, c.y_uq, c.cb_uq, c.cr_uq)
    for g, wv, name in zip(got, want, ("y_uq", "cb_uq", "cr_uq")):
        assert np.array_equal(g, wv), f"{name}: {np.count_nonzero(g != wv)} values differ"
    # quantising the side output on the host gives the quantised planes back (jcdctmgr.c:408-438)
    q = oracle.quantize(got[0][0, 0].astype(np.int32), np.array(ql, np.uint16))
    assert np.array_equal(np.asarray(q).reshape(-1), _coef_planes(c)[0][0, 0])


def test_jpeg_unquantised_batch_device():
    import torch
    n, w, h = 3, 130, 70
    imgs = np.stack([noise_image(w, h, 66 + i) for i in range(n)])
    src = torch.from_numpy(imgs).cuda()
    g = ll.jpeg_geometry(w, h)
    ql, qc = ll.jpeg_quant_table(70, 0), ll.jpeg_quant_table(70, 1)
    mk = lambda hb, wb: torch.zeros((n, hb, wb, 64), dtype=torch.int16, device="cuda")
    y, cb, cr, yu, cbu, cru = mk(g.y_HB, g.y_WB), mk(g.c_HB, g.c_WB), mk(g.c_HB, g.c_WB), mk(g.y_HB, g.y_WB), mk(g.c_HB, g.c_WB), mk(g.c_HB, g.c_WB)
    ll.jpeg_frontend_batch_device_ex(src.data_ptr(), n, w, h, ql, qc, y.data_ptr(), cb.data_ptr(), cr.data_ptr(),
                                     yu.data_ptr(), cbu.data_ptr(), cru.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for i in range(n):
        wq = oracle.jpeg_frontend(imgs[i], np.array(ql, np.uint16), np.array(qc, np.uint16))
        wu = oracle.jpeg_frontend_unquantised(imgs[i])
        for t, wv in zip((y, cb, cr, yu, cbu, cru), wq + wu):
            assert np.array_equal(t[i].cpu().numpy(), wv)


def test_pipeline_jpeg_batch_device_keeps_pixels_in_hbm():
    import torch
    n, w, h = 4, 600, 400
    imgs = np.stack([noise_image(w, h, 70 + i) for i in range(n)])
    src = torch.from_numpy(imgs).cuda()
    _, l = _modes(("width", 250), ("aspect", 16, 9), ("aspect", 1, 1), WHITE)
    ow, oh = ll.pipeline_output_dims(w, h, l, ll.ORDER_CHAIN)
    g = ll.jpeg_geometry(ow, oh)
    ql, qc = ll.jpeg_quant_table(70, 0), ll.jpeg_quant_table(70, 1)
    px = torch.zeros((n, oh, ow, 4), dtype=torch.uint8, device="cuda")
    y = torch.zeros((n, g.y_HB, g.y_WB, 64), dtype=torch.int16, device="cuda")
    cb = torch.zeros((n, g.c_HB, g.c_WB, 64), dtype=torch.int16, device="cuda")
    cr = torch.zeros_like(cb)
    before = ll.kernel_launches()
    ll.pipeline_jpeg_batch_device(src.data_ptr(), n, w, h, l, ll.ORDER_CHAIN, ql, qc, px.data_ptr(), px.numel(),
                                  y.data_ptr(), cb.data_ptr(), cr.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert ll.kernel_launches() == before + 2
    for i in range(n):
        want_px = _np(ll.pipeline(_ll_img(imgs[i]), l, ll.ORDER_CHAIN))
        assert np.array_equal(px[i].cpu().numpy(), want_px)
        wq = oracle.jpeg_frontend(want_px, np.array(ql, np.uint16), np.array(qc, np.uint16))
        for t, wv in zip((y, cb, cr), wq):
            assert np.array_equal(t[i].cpu().numpy(), wv)


# ---- hand-off to the reference library's entropy coder (product side: slimg_b200/jpeg_handoff) -------------------

def _handoff():
    from slimg_b200 import _host_codecs as hc
    if hc.handoff_library() is None:
        pytest.skip("libslimg_jpeg_handoff.so not built")
    return hc


@pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")
def test_convert_jpeg_file_is_byte_identical_to_the_cpu_encoder(monkeypatch):
    """convert(img, JPEG q70, crop + extend [+ resize]) with the front end on the GPU and libjpeg-turbo's entropy coder
    behind jpeg_write_coefficients (jctrans.c:43): the FILE equals what the full CPU encoder (colour conversion,
    downsampling, DCT, quantisation, Huffman -- all on the host) writes for the same pixels."""
    hc = _handoff()
    monkeypatch.setattr(ll, "JPEG_TABLES", ll.QT_ANNEX_K)                    # libjpeg-turbo's own tables
    hc.install()                                                             # then replace the entropy coder's settings:
    ll.register_jpeg_entropy_coder(lambda coefs, q: hc.jpeg_from_coefficients(coefs, hc.HANDOFF_OPTIMIZE))
    try:
        img = noise_image(401, 299, 71)
        ql, qc = oracle.ref_jpeg_quant_table(70, 0), oracle.ref_jpeg_quant_table(70, 1)
        # crop + extend: pixels are bit-exact, so the file is pinned to the ORACLE's pixels
        opts = ll.PipelineOptions(format=ll.Format.JPEG, quality=70, crop=ll.CropMode.REGION(7, 9, 300, 200),
                                  extend=ll.ExtendMode.SIZE(333, 257), fill_color=ll.FillColor.SOLID(10, 200, 30, 128))
        res = ll.convert(_ll_img(img), opts)
        want_px = oracle.pipeline(img, crop_mode=oracle.CropMode.region(7, 9, 300, 200),
                                  extend_mode=oracle.ExtendMode.size(333, 257), fill=(10, 200, 30, 128))
        assert (res.width, res.height) == (333, 257) and res.data == oracle.ref_jpeg_encode(want_px, ql, qc)
        # with a resize: pinned to the pixels the GPU pipeline returns (<= 1 LSB from the oracle's on rounding ties)
        opts = ll.PipelineOptions(format=ll.Format.JPEG, quality=70, resize=ll.ResizeMode.WIDTH(160),
                                  crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1))
        res = ll.convert(_ll_img(img), opts)
        px = _np(ll.pipeline(_ll_img(img), opts._to_c(), ll.ORDER_CONVERT))
        assert res.data == oracle.ref_jpeg_encode(px, ql, qc)
        _close(px, oracle.pipeline(img, crop_mode=oracle.CropMode.aspect_ratio(16, 9), extend_mode=oracle.ExtendMode.aspect_ratio(1, 1),
                                   fill=WHITE, resize_mode=oracle.ResizeMode.width(160)), "noise", "convert jpeg pixels")
        # the reference's own settings (progressive + optimised tables, codec/jpeg.rs:61-63): a valid JPEG of that size
        prog = hc.jpeg_from_coefficients(ll.pipeline_jpeg(_ll_img(img), opts._to_c(), list(ql), list(qc)))
        assert prog[:2] == bytes([0xff, 0xd8]) and oracle.ref_jpeg_decode(prog, 160, 160).shape == (160, 160, 4)
        assert np.array_equal(oracle.ref_jpeg_decode(prog, 160, 160), oracle.ref_jpeg_decode(res.data, 160, 160))
    finally:
        ll.register_jpeg_entropy_coder(lambda coefs, q: hc.jpeg_from_coefficients(coefs))      # back to the reference's settings


# ---- batches: mixed sizes, pageable and pinned buffers, plan cache churn -------------------------------------------------

def _batch(imgs, opts, order, pinned, cfg=None):
    lib = ll.library()
    n = len(imgs)
    cin = (ll.CImage * n)()
    keep, pins = [], []
    for i, a in enumerate(imgs):
        raw = a.tobytes()
        if pinned:
            p = lib.slimg_b200_host_alloc(len(raw))
            C.memmove(p, raw, len(raw))
            pins.append(p)
            cin[i] = ll.CImage(a.shape[1], a.shape[0], p, len(raw))
        else:
            keep.append(raw)
            cin[i] = ll.CImage(a.shape[1], a.shape[0], C.cast(C.c_char_p(raw), C.c_void_p), len(raw))
    cout = (ll.CImage * n)()
    st = (C.c_int * n)()
    failed = lib.slimg_b200_pipeline_batch_host(cin, n, C.byref(opts), order, cfg, cout, st)
    outs = []
    for i in range(n):
        if st[i] == 0:
            outs.append(np.frombuffer(C.string_at(cout[i].data, cout[i].len), np.uint8).reshape(cout[i].height, cout[i].width, 4).copy())
            lib.slimg_b200_image_free(C.byref(cout[i]))
        else:
            outs.append(None)
    for p in pins:
        lib.slimg_b200_host_free(p)
    return failed, list(st), outs


def test_batch_host_directory_like_mix_pageable_equals_pinned():
    """What --jobs sees (cli/src/commands/convert.rs:46-60): photos of different sizes and aspect ratios in one batch,
    out of ordinary heap memory.  Every image equals its single-image result; pageable staging == pinned."""
    sizes = [(4000, 3000), (3000, 4000), (1024, 1024), (1920, 1080), (2048, 1365), (1000, 334), (4000, 3000), (640, 481)]
    imgs = [noise_image(w, h, 80 + i) for i, (w, h) in enumerate(sizes)]
    opts = ll.make_options(resize=ll.ResizeMode.FIT(400, 400), crop=ll.CropMode.ASPECT_RATIO(4, 3))
    f1, st1, pageable = _batch(imgs, opts, ll.ORDER_CONVERT, pinned=False)
    f2, st2, pinned = _batch(imgs, opts, ll.ORDER_CONVERT, pinned=True)
    assert f1 == 0 and f2 == 0 and st1 == st2
    for a, b, img in zip(pageable, pinned, imgs):
        assert np.array_equal(a, b)
        assert np.array_equal(a, _np(ll.pipeline(_ll_img(img), opts, ll.ORDER_CONVERT)))
    for i in (2, 5, 7):          # the small ones against the oracle (the big ones are covered by the full-size cases)
        want = oracle.pipeline(imgs[i], crop_mode=oracle.CropMode.aspect_ratio(4, 3), resize_mode=oracle.ResizeMode.fit(400, 400))
        _close(pageable[i], want, "noise", f"mixed batch {sizes[i]}")


def test_batch_host_pageable_large_copy_through_the_staging_ring():
    """A source larger than both staging slots (2 x 8 MiB) and an output larger than one: chunked both ways."""
    img = noise_image(4000, 3000, 90)                 # 48 MB in, 36 MB out
    opts = ll.make_options(crop=ll.CropMode.ASPECT_RATIO(16, 9))
    failed, st, outs = _batch([img, img], opts, ll.ORDER_CONVERT, pinned=False)
    assert failed == 0
    want = oracle.crop(img, oracle.CropMode.aspect_ratio(16, 9))
    assert np.array_equal(outs[0], want) and np.array_equal(outs[1], want)


def test_plan_cache_survives_more_geometries_than_it_holds():
    """A corpus in which every photo has its own size: 80 geometries through a 64-entry LRU, two passes."""
    opts = ll.make_options(resize=ll.ResizeMode.SCALE(0.5))
    for rep in range(2):
        for k in range(80):
            w, h = 64 + 4 * k, 48 + 2 * (k % 7)
            img = noise_image(w, h, k)
            got = _np(ll.pipeline(_ll_img(img), opts, ll.ORDER_CONVERT))
            if rep == 1 and k % 9 == 0:
                _close(got, oracle.resize(img, oracle.ResizeMode.scale(0.5)), "noise", f"plan cache churn {w}x{h}")


def test_unique_geometries_through_the_scheduler_recycle_plan_slabs():
    """Every image its own geometry, eight workers building and dropping plans at once (api.cu: SlabPool -- retired slabs
    are recycled behind one device synchronise per 16, never cudaFree'd on the hot path).  300 geometries through the
    64-entry cache, three batches; every output equals the single-image call's."""
    opts = ll.make_options(resize=ll.ResizeMode.SCALE(0.5))
    dev = (C.c_int32 * 1)(0)
    cfg = ll.CBatchConfig(1, 8, dev)
    for rep in range(3):
        sizes = [(96 + 4 * ((rep * 100 + k) % 300), 40 + (k % 5)) for k in range(100)]
        imgs = [noise_image(w, h, (rep * 100 + k) % 50) for k, (w, h) in enumerate(sizes)]
        failed, st, outs = _batch(imgs, opts, ll.ORDER_CONVERT, pinned=False, cfg=C.byref(cfg))
        assert failed == 0 and not any(st)
        for k in range(0, 100, 7):
            assert np.array_equal(outs[k], _np(ll.pipeline(_ll_img(imgs[k]), opts, ll.ORDER_CONVERT))), (rep, k)


def test_batch_jpeg_host_scheduler():
    lib = ll.library()
    sizes = [(640, 480), (333, 257), (3, 3), (1000, 334)]
    imgs = [noise_image(w, h, 95 + i) for i, (w, h) in enumerate(sizes)]
    keep = [a.tobytes() for a in imgs]
    n = len(imgs)
    cin = (ll.CImage * n)()
    for i, a in enumerate(imgs):
        cin[i] = ll.CImage(a.shape[1], a.shape[0], C.cast(C.c_char_p(keep[i]), C.c_void_p), len(keep[i]))
    cout = (ll.CJpegCoefs * n)()
    st = (C.c_int * n)()
    ql, qc = ll.jpeg_quant_table(70, 0), ll.jpeg_quant_table(70, 1)
    qlc, qcc = (C.c_uint16 * 64)(*ql), (C.c_uint16 * 64)(*qc)
    opts = ll.make_options(resize=ll.ResizeMode.SCALE(0.5), crop=ll.CropMode.REGION(0, 0, 10, 10))
    failed = lib.slimg_b200_pipeline_jpeg_batch_host(cin, n, C.byref(opts), ll.ORDER_CONVERT, qlc, qcc, 0, None, cout, st)
    assert failed == 1 and st[2] == 7                                   # the 3x3 image cannot be cropped to 10x10
    for i in (0, 1, 3):
        want_px = _np(ll.pipeline(_ll_img(imgs[i]), opts, ll.ORDER_CONVERT))
        want = oracle.jpeg_frontend(want_px, np.array(ql, np.uint16), np.array(qc, np.uint16))
        g = cout[i].geom
        got_y = np.frombuffer(C.string_at(cout[i].y, g.y_HB * g.y_WB * 128), np.int16).reshape(g.y_HB, g.y_WB, 64)
        assert (g.width, g.height) == (want_px.shape[1], want_px.shape[0]) and np.array_equal(got_y, want[0])
        lib.slimg_b200_jpeg_coefs_free(C.byref(cout[i]))


def test_in_library_multi_device_scheduler_uses_every_visible_gpu():
    """slimg_b200_pipeline_batch_host with n_devices = all: the --jobs analogue inside ONE process."""
    n_dev = ll.device_count()
    imgs = [noise_image(800, 600, 100 + i) for i in range(4 * max(n_dev, 1))]
    opts = ll.make_options(resize=ll.ResizeMode.WIDTH(200))
    cfg = ll.CBatchConfig(n_dev, 2, None)
    failed, st, outs = _batch(imgs, opts, ll.ORDER_CONVERT, pinned=False, cfg=C.byref(cfg))
    assert failed == 0
    for img, out in zip(imgs, outs):
        assert np.array_equal(out, _np(ll.pipeline(_ll_img(img), opts, ll.ORDER_CONVERT)))


# ---- MozJPEG's front-end defaults: overshoot deringing + 10-bit clamp (restated, unpinned) -----------------------------

def _line_art(w, h, seed):
    """White paper, dark strokes: the clipped whites MozJPEG's deringing is made for."""
    rng = np.random.default_rng(seed)
    img = np.full((h, w, 4), 255, np.uint8)
    for _ in range(max(4, w * h // 900)):
        x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
        img[y:y + int(rng.integers(1, 9)), x:x + int(rng.integers(1, 40)), :3] = rng.integers(0, 200, 3)
    return img


@pytest.mark.parametrize("kind,w,h,q", [("lineart", 640, 480, 70), ("lineart", 333, 257, 30), ("noise", 250, 131, 70), ("white", 64, 48, 70),
                                        ("gradient", 320, 200, 90), ("lineart", 1920, 1080, 100)])
def test_jpeg_frontend_with_mozjpeg_deringing(kind, w, h, q):
    img = {"lineart": lambda: _line_art(w, h, 5), "noise": lambda: noise_image(w, h, 12), "white": lambda: constant_image(w, h, 255),
           "gradient": lambda: gradient_image(w, h)}[kind]()
    ql, qc = ll.jpeg_quant_table(q, 0, ll.QT_MOZJPEG), ll.jpeg_quant_table(q, 1, ll.QT_MOZJPEG)
    want = oracle.jpeg_frontend_mozjpeg(img, np.array(ql, np.uint16), np.array(qc, np.uint16))
    plain = oracle.jpeg_frontend(img, np.array(ql, np.uint16), np.array(qc, np.uint16))
    c = ll.jpeg_frontend_ex(_ll_img(img), ql, qc, ll.JPEG_DERINGING)
    for g, wv, name in zip(_coef_planes(c), want, ("y", "cb", "cr")):
        assert np.array_equal(g, wv), f"{name}: {np.count_nonzero(g != wv)} coefficients differ"
    if kind == "lineart":
        assert any(not np.array_equal(a, b) for a, b in zip(want, plain))       # the stage did something
    if kind in ("white", "gradient") and q < 100:
        assert all(np.array_equal(a, b) for a, b in zip(want, plain))           # nothing to dering, nothing clamped
    # the side output is the DCT of the deringed block (what MozJPEG keeps for trellis)
    cu = ll.jpeg_frontend_ex(_ll_img(img), ql, qc, ll.JPEG_DERINGING | ll.JPEG_UNQUANTISED)
    assert (cu.y, cu.cb, cu.cr) == (c.y, c.cb, c.cr) and cu.y_uq is not None
