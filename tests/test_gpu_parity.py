"""Parity of the CUDA path (through the C ABI) with the oracle on seeded inputs.
Bars (BASELINE.json north_star): crop / extend bit-exact; resize within +-1 LSB per u8 channel with
the mean abs error reported and bounded; JPEG coefficients bit-exact (integer pipeline)."""
import ctypes as C

import numpy as np
import pytest

import oracle
from conftest import IMAGE_KINDS, binding_image, gradient_image, noise_image
from golden_util import jpeg_cases, resize_cases
from slimg_b200 import _lowlevel as ll

pytestmark = pytest.mark.gpu

WHITE, CLEAR = (255, 255, 255, 255), (0, 0, 0, 0)


def _ll_img(a):
    return ll.ImageData(a.shape[1], a.shape[0], a.tobytes())


def _np(img):
    return np.frombuffer(img.data, np.uint8).reshape(img.height, img.width, 4)


def _fill_ll(f):
    return ll.FillColor.TRANSPARENT() if f == CLEAR else ll.FillColor.SOLID(*f)


# Per-kind bounds on the MEAN abs error (the max is always <= 1 LSB, BASELINE.json).  noise / alpha_ramp: a correct
# kernel flips well under 1 % of the roundings.  constant: the fixed-point weights sum exactly (vplan.cpp: quantise), so
# a constant image must come back exactly.  gradient / checker at exact ratios put many samples EXACTLY on .5, where the
# reference's own f32 rounding noise picks the side and any other summation order disagrees on part of them: bounds
# are twice the largest value measured over all cases of that kind (profiles/r02_parity.txt).
MEAN_BOUND = {"noise": 0.02, "alpha_ramp": 0.02, "constant": 0.0, "gradient": 0.30, "checker": 0.10, "guard": 0.05}
from conftest import PARITY_REPORT as _REPORT        # written to gpurun_out/parity_report.txt at session end


def _close(got, want, kind="noise", label=""):
    assert got.shape == want.shape, (got.shape, want.shape)
    d = np.abs(got.astype(np.int16) - want.astype(np.int16))
    mx, mean = int(d.max()) if d.size else 0, float(d.mean()) if d.size else 0.0
    _REPORT.append((label or kind, kind, got.shape[1], got.shape[0], mx, mean, ll.PATH_NAMES[ll.last_resize_path()]))
    assert mx <= 1, f"max abs err {mx}"
    assert mean <= MEAN_BOUND[kind], f"mean abs err {mean} (bound {MEAN_BOUND[kind]} for {kind})"
    return mx, mean


# ---- K1: crop / extend, bit-exact ---------------------------------------------------------------

@pytest.mark.parametrize("w,h,mode", [
    (10, 8, ("region", 2, 1, 5, 4)), (200, 100, ("aspect", 1, 1)), (100, 100, ("aspect", 16, 9)),
    (4001, 2999, ("region", 3, 7, 3990, 2900)), (1000, 334, ("aspect", 4, 3)), (17, 13, ("region", 16, 12, 1, 1)),
    (64, 64, ("region", 0, 0, 64, 64)), (1023, 5, ("region", 1, 0, 1020, 5)),
])
def test_crop_bit_exact(w, h, mode):
    img = noise_image(w, h)
    om = oracle.CropMode.region(*mode[1:]) if mode[0] == "region" else oracle.CropMode.aspect_ratio(*mode[1:])
    lm = ll.CropMode.REGION(*mode[1:]) if mode[0] == "region" else ll.CropMode.ASPECT_RATIO(*mode[1:])
    assert np.array_equal(_np(ll.crop(_ll_img(img), lm)), oracle.crop(img, om))


@pytest.mark.parametrize("w,h,mode,fill", [
    (10, 8, ("aspect", 1, 1), CLEAR), (6, 10, ("aspect", 1, 1), CLEAR), (4, 4, ("size", 6, 6), (255, 0, 0, 255)),
    (4, 4, ("size", 6, 6), (128, 64, 32, 200)), (200, 100, ("aspect", 2, 1), WHITE), (4, 4, ("size", 7, 9), (9, 8, 7, 6)),
    (1000, 334, ("size", 1003, 1001), (1, 2, 3, 4)), (4000, 2250, ("aspect", 1, 1), CLEAR), (17, 13, ("size", 17, 400), WHITE),
    (2, 1, ("size", 4, 3), CLEAR),
])
def test_extend_bit_exact(w, h, mode, fill):
    img = noise_image(w, h, 1)
    om = oracle.ExtendMode.aspect_ratio(*mode[1:]) if mode[0] == "aspect" else oracle.ExtendMode.size(*mode[1:])
    lm = ll.ExtendMode.ASPECT_RATIO(*mode[1:]) if mode[0] == "aspect" else ll.ExtendMode.SIZE(*mode[1:])
    assert np.array_equal(_np(ll.extend(_ll_img(img), lm, _fill_ll(fill))), oracle.extend(img, om, fill))


def test_config2_crop_16_9_extend_1_1_transparent_4000x3000():
    """BASELINE config 2, one fused launch, bit-exact."""
    img = noise_image(4000, 3000, 2)
    opts = ll.make_options(crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1),
                           fill_color=ll.FillColor.TRANSPARENT())
    got = _np(ll.pipeline(_ll_img(img), opts, ll.ORDER_CONVERT))
    want = oracle.pipeline(img, crop_mode=oracle.CropMode.aspect_ratio(16, 9),
                           extend_mode=oracle.ExtendMode.aspect_ratio(1, 1), fill=CLEAR)
    assert got.shape == (4000, 4000, 4) and np.array_equal(got, want)
    # same thing as two standalone calls (SlimgTest.kt:324-341)
    two = ll.extend(ll.crop(_ll_img(img), ll.CropMode.ASPECT_RATIO(16, 9)), ll.ExtendMode.ASPECT_RATIO(1, 1),
                    ll.FillColor.TRANSPARENT())
    assert np.array_equal(_np(two), want)


def test_default_fill_is_opaque_white():             # pipeline.rs:78-80
    img = noise_image(40, 30, 3)
    got = _np(ll.pipeline(_ll_img(img), ll.make_options(extend=ll.ExtendMode.SIZE(50, 50)), ll.ORDER_CONVERT))
    assert np.array_equal(got, oracle.extend(img, oracle.ExtendMode.size(50, 50), WHITE))


# ---- K2: Lanczos3 resize, +-1 LSB ------------------------------------------------------------------

RESIZE_CASES = [
    ("noise", 400, 300, ("width", 80)), ("noise", 300, 400, ("height", 77)), ("gradient", 512, 512, ("width", 256)),
    ("noise", 64, 64, ("scale", 4.0)), ("checker", 53, 37, ("exact", 17, 41)), ("alpha_ramp", 200, 100, ("fit", 64, 64)),
    ("noise", 1000, 334, ("width", 300)), ("noise", 10, 8, ("scale", 2.0)), ("noise", 10, 8, ("width", 5)),
    ("constant", 200, 100, ("exact", 50, 50)), ("noise", 17, 13, ("exact", 1, 1)), ("noise", 1, 1, ("exact", 9, 5)),
    ("noise", 2000, 31, ("exact", 40, 90)), ("noise", 257, 129, ("scale", 0.37)), ("checker", 128, 128, ("scale", 1.5)),
    ("noise", 900, 700, ("exact", 9, 7)),
]


@pytest.mark.parametrize("kind,w,h,mode", RESIZE_CASES, ids=lambda v: str(v))
def test_resize_within_one_lsb(kind, w, h, mode):
    img = IMAGE_KINDS[kind](w, h)
    want = oracle.resize(img, getattr(oracle.ResizeMode, mode[0])(*mode[1:]))
    got = _np(ll.resize(_ll_img(img), getattr(ll.ResizeMode, mode[0].upper())(*mode[1:])))
    _close(got, want, kind, f"resize {kind} {w}x{h} {mode}")


@pytest.mark.parametrize("case", list(resize_cases()), ids=lambda c: c["id"])
def test_resize_golden_vectors(case):
    got = _np(ll.resize(_ll_img(case["img"]), ll.ResizeMode.EXACT(case["nw"], case["nh"])))
    _close(got, case["out"], case["id"].split("-")[0] if case["id"].split("-")[0] in MEAN_BOUND else "noise", "golden " + case["id"])


def test_resize_same_size_is_a_copy():
    img = noise_image(33, 21)
    assert np.array_equal(_np(ll.resize(_ll_img(img), ll.ResizeMode.WIDTH(33))), img)


PIPE_CASES = [
    # (w, h, resize, crop, extend, fill)
    (600, 400, ("width", 80), ("aspect", 16, 9), ("aspect", 1, 1), WHITE),
    (600, 400, ("width", 80), ("aspect", 16, 9), ("aspect", 1, 1), CLEAR),
    (401, 299, ("scale", 0.25), ("region", 7, 9, 300, 200), ("size", 333, 257), (10, 200, 30, 128)),
    (120, 90, ("scale", 3.0), ("aspect", 1, 1), ("aspect", 16, 9), (0, 0, 0, 255)),
    (300, 200, ("fit", 100, 100), None, ("aspect", 1, 1), None),
    (300, 200, ("height", 50), ("region", 10, 10, 100, 150), None, None),
]


def _modes(resize, crop, extend, fill):
    o = dict(resize_mode=getattr(oracle.ResizeMode, resize[0])(*resize[1:]) if resize else None,
             crop_mode=(oracle.CropMode.region(*crop[1:]) if crop[0] == "region" else oracle.CropMode.aspect_ratio(*crop[1:])) if crop else None,
             extend_mode=(oracle.ExtendMode.aspect_ratio(*extend[1:]) if extend[0] == "aspect" else oracle.ExtendMode.size(*extend[1:])) if extend else None,
             fill=fill)
    l = ll.make_options(
        resize=getattr(ll.ResizeMode, resize[0].upper())(*resize[1:]) if resize else None,
        crop=(ll.CropMode.REGION(*crop[1:]) if crop[0] == "region" else ll.CropMode.ASPECT_RATIO(*crop[1:])) if crop else None,
        extend=(ll.ExtendMode.ASPECT_RATIO(*extend[1:]) if extend[0] == "aspect" else ll.ExtendMode.SIZE(*extend[1:])) if extend else None,
        fill_color=_fill_ll(fill) if fill is not None else None)
    return o, l


@pytest.mark.parametrize("w,h,resize,crop,extend,fill", PIPE_CASES)
def test_convert_order_fill_takes_part_in_the_taps(w, h, resize, crop, extend, fill):
    img = noise_image(w, h, 5)
    o, l = _modes(resize, crop, extend, fill)
    want = oracle.pipeline(img, **o)                                  # crop -> extend -> resize
    _close(_np(ll.pipeline(_ll_img(img), l, ll.ORDER_CONVERT)), want, "noise", f"convert-order {w}x{h} {resize}")


CHAIN_CASES = [
    (600, 400, ("width", 80), ("aspect", 16, 9), ("aspect", 1, 1), WHITE),
    (600, 400, ("width", 80), ("aspect", 16, 9), ("aspect", 1, 1), CLEAR),
    (401, 299, ("scale", 0.25), ("region", 7, 9, 80, 50), ("size", 133, 57), (10, 200, 30, 128)),
    (120, 90, ("scale", 3.0), ("aspect", 1, 1), ("aspect", 16, 9), (0, 0, 0, 255)),
    (300, 200, ("fit", 100, 100), None, ("aspect", 1, 1), None),
    (300, 200, ("height", 50), ("region", 10, 10, 60, 35), None, None),
    (1203, 807, ("width", 301), ("region", 1, 2, 299, 198), ("size", 640, 480), (1, 2, 3, 255)),
]


@pytest.mark.parametrize("w,h,resize,crop,extend,fill", CHAIN_CASES)
def test_chain_order_resize_crop_extend(w, h, resize, crop, extend, fill):
    img = noise_image(w, h, 6)
    o, l = _modes(resize, crop, extend, fill)
    want = oracle.resize(img, o["resize_mode"])
    if o["crop_mode"] is not None:
        want = oracle.crop(want, o["crop_mode"])
    if o["extend_mode"] is not None:
        want = oracle.extend(want, o["extend_mode"], WHITE if fill is None else fill)
    got = _np(ll.pipeline(_ll_img(img), l, ll.ORDER_CHAIN))
    _close(got, want, "noise", f"chain-order {w}x{h} {resize}")
    # the padding itself must be bit-exact
    if o["extend_mode"] is not None:
        pad = np.ones(want.shape[:2], bool)
        ph, pw = oracle.crop(oracle.resize(img, o["resize_mode"]), o["crop_mode"]).shape[:2] if o["crop_mode"] is not None \
            else oracle.resize(img, o["resize_mode"]).shape[:2]
        oy, ox = (want.shape[0] - ph) // 2, (want.shape[1] - pw) // 2
        pad[oy:oy + ph, ox:ox + pw] = False
        assert np.array_equal(got[pad], want[pad])


@pytest.mark.parametrize("name,w,h,mode", [
    ("C1", 4000, 3000, ("width", 800)), ("C4a", 7680, 4320, ("width", 1920)), ("C4b", 1024, 1024, ("scale", 4.0)),
])
def test_baseline_resize_configs_full_size(name, w, h, mode):
    img = noise_image(w, h, 8)
    want = oracle.resize(img, getattr(oracle.ResizeMode, mode[0])(*mode[1:]))
    got = _np(ll.resize(_ll_img(img), getattr(ll.ResizeMode, mode[0].upper())(*mode[1:])))
    mx, mean = _close(got, want, "noise", f"{name} full size")
    print(f"{name}: max abs err {mx}, mean abs err {mean:.6f}")


@pytest.mark.parametrize("order", ["convert", "chain"])
def test_config3_24mp_full_size(order):
    """BASELINE config 3 on one 6000x4000 image, both stage orders."""
    img = noise_image(6000, 4000, 9)
    o, l = _modes(("width", 800), ("aspect", 16, 9), ("aspect", 1, 1), WHITE)
    if order == "convert":
        want = oracle.pipeline(img, **o)
        got = _np(ll.pipeline(_ll_img(img), l, ll.ORDER_CONVERT))
    else:
        want = oracle.extend(oracle.crop(oracle.resize(img, o["resize_mode"]), o["crop_mode"]), o["extend_mode"], WHITE)
        got = _np(ll.pipeline(_ll_img(img), l, ll.ORDER_CHAIN))
    assert got.shape == (800, 800, 4)
    mx, mean = _close(got, want, "noise", f"C3 {order} full size")
    print(f"C3 {order}: max abs err {mx}, mean abs err {mean:.6f}")


# ---- K3/K4: JPEG front end, bit-exact ---------------------------------------------------------------

def _planes(geom, y, cb, cr):
    return (np.frombuffer(y, np.int16).reshape(geom.y_HB, geom.y_WB, 64),
            np.frombuffer(cb, np.int16).reshape(geom.c_HB, geom.c_WB, 64),
            np.frombuffer(cr, np.int16).reshape(geom.c_HB, geom.c_WB, 64))


@pytest.mark.parametrize("case", list(jpeg_cases()), ids=lambda c: c["id"])
def test_jpeg_golden_from_real_libjpeg_turbo(case):
    got = _planes(*ll.jpeg_frontend(_ll_img(case["img"]), list(case["ql"]), list(case["qc"])))
    for g, name in zip(got, ("y", "cb", "cr")):
        assert np.array_equal(g, case[name]), name


@pytest.mark.parametrize("w,h,q,base", [(16, 16, 70, 0), (17, 13, 70, 1), (129, 65, 30, 0), (640, 481, 70, 1),
                                        (641, 480, 95, 0), (250, 250, 100, 0), (1, 1, 70, 0), (2000, 9, 70, 0),
                                        (6000, 4000, 70, 1)])
def test_jpeg_frontend_bit_exact(w, h, q, base):
    img = noise_image(w, h, 11)
    ql, qc = oracle.jpeg_quant_table(q, 0, base), oracle.jpeg_quant_table(q, 1, base)
    want = oracle.jpeg_frontend(img, ql, qc)
    got = _planes(*ll.jpeg_frontend(_ll_img(img), list(ql), list(qc)))
    for g, wv, name in zip(got, want, ("y", "cb", "cr")):
        assert g.shape == wv.shape
        assert np.array_equal(g, wv), f"{name}: {np.count_nonzero(g != wv)} coefficients differ"


@pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")
@pytest.mark.parametrize("w,h", [(53, 37), (640, 481), (1920, 1080)])
def test_jpeg_coefficients_feed_the_reference_entropy_coder(w, h):
    """the GPU front end's coefficients go to libjpeg-turbo's entropy coder through
    jpeg_write_coefficients (jctrans.c:43); the JPEG file is byte-identical to the one the full CPU encoder writes."""
    img = noise_image(w, h, 17)
    ql, qc = oracle.ref_jpeg_quant_table(70, 0), oracle.ref_jpeg_quant_table(70, 1)
    y, cb, cr = _planes(*ll.jpeg_frontend(_ll_img(img), list(ql), list(qc)))
    got = oracle.ref_jpeg_write_coefficients(y, cb, cr, w, h, ql, qc)
    assert got == oracle.ref_jpeg_encode(img, ql, qc)
    prog = oracle.ref_jpeg_write_coefficients(y, cb, cr, w, h, ql, qc, progressive=True)       # codec/jpeg.rs:61
    assert oracle.ref_jpeg_decode(prog, w, h).shape == (h, w, 4)


# ---- batched device-resident entry points (what bench.py times) ---------------------------------------

def test_batch_device_matches_single_image_calls():
    import torch
    n, w, h = 5, 300, 200
    imgs = np.stack([noise_image(w, h, 20 + i) for i in range(n)])
    src = torch.from_numpy(imgs).cuda()
    for order in (ll.ORDER_CONVERT, ll.ORDER_CHAIN):
        _, l = _modes(("width", 64), ("aspect", 16, 9), ("aspect", 1, 1), (1, 2, 3, 4))
        ow, oh = ll.pipeline_output_dims(w, h, l, order)
        dst = torch.zeros((n, oh, ow, 4), dtype=torch.uint8, device="cuda")
        before = ll.kernel_launches()
        ll.pipeline_batch_device(src.data_ptr(), n, w, h, l, order, dst.data_ptr(), dst.numel(),
                                 torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert ll.kernel_launches() == before + 1
        out = dst.cpu().numpy()
        for i in range(n):
            assert np.array_equal(out[i], _np(ll.pipeline(_ll_img(imgs[i]), l, order)))
    with pytest.raises(ll.SlimgError.Image):
        ll.pipeline_batch_device(src.data_ptr(), n, w, h, l, 0, dst.data_ptr(), 16, 0)


def test_jpeg_batch_device():
    import torch
    n, w, h = 3, 100, 60
    imgs = np.stack([noise_image(w, h, 30 + i) for i in range(n)])
    src = torch.from_numpy(imgs).cuda()
    g = ll.jpeg_geometry(w, h)
    ql, qc = ll.jpeg_quant_table(70, 0), ll.jpeg_quant_table(70, 1)
    y = torch.zeros((n, g.y_HB, g.y_WB, 64), dtype=torch.int16, device="cuda")
    cb = torch.zeros((n, g.c_HB, g.c_WB, 64), dtype=torch.int16, device="cuda")
    cr = torch.zeros_like(cb)
    ll.jpeg_frontend_batch_device(src.data_ptr(), n, w, h, ql, qc, y.data_ptr(), cb.data_ptr(), cr.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for i in range(n):
        want = oracle.jpeg_frontend(imgs[i], np.array(ql, np.uint16), np.array(qc, np.uint16))
        assert np.array_equal(y[i].cpu().numpy(), want[0])
        assert np.array_equal(cb[i].cpu().numpy(), want[1])
        assert np.array_equal(cr[i].cpu().numpy(), want[2])


def test_batch_host_scheduler_isolates_failures():
    """--jobs analogue: mixed sizes, one image that cannot be cropped; the others still complete
    (ErrorCollector semantics, cli/src/commands/mod.rs:113-145)."""
    lib = ll.library()
    sizes = [(64, 48), (100, 100), (3, 3), (200, 120), (31, 77)]
    imgs = [noise_image(w, h, 40 + i) for i, (w, h) in enumerate(sizes)]
    keep = [a.tobytes() for a in imgs]
    cin = (ll.CImage * len(imgs))()
    for i, a in enumerate(imgs):
        cin[i] = ll.CImage(a.shape[1], a.shape[0], C.cast(C.c_char_p(keep[i]), C.c_void_p), len(keep[i]))
    cout = (ll.CImage * len(imgs))()
    st = (C.c_int * len(imgs))()
    opts = ll.make_options(resize=ll.ResizeMode.SCALE(0.5), crop=ll.CropMode.REGION(0, 0, 10, 10))
    failed = lib.slimg_b200_pipeline_batch_host(cin, len(imgs), C.byref(opts), ll.ORDER_CONVERT, None, cout, st)
    assert failed == 1 and st[2] == 7 and [st[i] for i in (0, 1, 3, 4)] == [0, 0, 0, 0]
    for i in (0, 1, 3, 4):
        got = np.frombuffer(C.string_at(cout[i].data, cout[i].len), np.uint8).reshape(cout[i].height, cout[i].width, 4)
        want = oracle.pipeline(imgs[i], crop_mode=oracle.CropMode.region(0, 0, 10, 10), resize_mode=oracle.ResizeMode.scale(0.5))
        _close(got, want, "noise", "batch_host mixed sizes")
        lib.slimg_b200_image_free(C.byref(cout[i]))


# ---- no kernel writes outside its output (the box has no compute-sanitizer: guard bands instead) -----------------

GUARD_CASES = [
    # (w, h, resize, crop, extend, fill, order)
    (1200, 800, ("width", 160), ("aspect", 16, 9), ("aspect", 1, 1), WHITE, "chain"),       # V-first + fill CTAs
    (1200, 800, ("width", 160), ("aspect", 16, 9), ("aspect", 1, 1), WHITE, "convert"),
    (150, 90, ("scale", 3.0), None, None, None, "convert"),                                   # few-tap kernel
    (333, 222, ("exact", 500, 100), None, None, None, "convert"),                             # generic kernel (up x, down y)
    (401, 299, None, ("region", 7, 9, 300, 200), ("size", 333, 257), (10, 200, 30, 128), "convert"),   # K1 remap
    (1001, 667, ("width", 125), None, ("aspect", 3, 2), CLEAR, "chain"),                      # odd pitch, unaligned rows
]


@pytest.mark.parametrize("w,h,resize,crop,extend,fill,order", GUARD_CASES)
def test_batch_device_writes_only_its_output(w, h, resize, crop, extend, fill, order):
    import torch
    n = 3
    imgs = np.stack([noise_image(w, h, 40 + i) for i in range(n)])
    src = torch.from_numpy(imgs).cuda()
    o, l = _modes(resize, crop, extend, fill)
    od = ll.ORDER_CHAIN if order == "chain" else ll.ORDER_CONVERT
    ow, oh = ll.pipeline_output_dims(w, h, l, od)
    nbytes, guard = n * oh * ow * 4, 1 << 16
    buf = torch.full((nbytes + 2 * guard,), 0xA5, dtype=torch.uint8, device="cuda")
    ll.pipeline_batch_device(src.data_ptr(), n, w, h, l, od, buf.data_ptr() + guard, nbytes,
                             torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    host = buf.cpu().numpy()
    assert (host[:guard] == 0xA5).all() and (host[guard + nbytes:] == 0xA5).all(), "kernel wrote outside its output"
    got = host[guard:guard + nbytes].reshape(n, oh, ow, 4)
    for i in range(n):
        if order == "convert":
            want = oracle.pipeline(imgs[i], **o)
        else:
            want = imgs[i]
            if o["resize_mode"] is not None: want = oracle.resize(want, o["resize_mode"])
            if o["crop_mode"] is not None: want = oracle.crop(want, o["crop_mode"])
            if o["extend_mode"] is not None: want = oracle.extend(want, o["extend_mode"], fill if fill is not None else WHITE)
        _close(got[i], want, "guard", f"guard band {w}x{h} {order}")


def test_jpeg_batch_device_writes_only_its_planes():
    import torch
    n, w, h = 2, 250, 131
    imgs = np.stack([noise_image(w, h, 50 + i) for i in range(n)])
    src = torch.from_numpy(imgs).cuda()
    g = ll.jpeg_geometry(w, h)
    ql, qc = ll.jpeg_quant_table(70, 0), ll.jpeg_quant_table(70, 1)
    ny, nc, guard = n * g.y_HB * g.y_WB * 64, n * g.c_HB * g.c_WB * 64, 4096
    bufs = [torch.full((k + 2 * guard,), 0x5A5A, dtype=torch.int16, device="cuda") for k in (ny, nc, nc)]
    ll.jpeg_frontend_batch_device(src.data_ptr(), n, w, h, ql, qc, *[b.data_ptr() + 2 * guard for b in bufs],
                                  torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for b, k in zip(bufs, (ny, nc, nc)):
        hb = b.cpu().numpy()
        assert (hb[:guard] == 0x5A5A).all() and (hb[guard + k:] == 0x5A5A).all(), "kernel wrote outside its plane"
    for i in range(n):
        want = oracle.jpeg_frontend(imgs[i], np.array(ql, np.uint16), np.array(qc, np.uint16))
        got_y = bufs[0].cpu().numpy()[guard:guard + ny].reshape(n, g.y_HB, g.y_WB, 64)[i]
        assert np.array_equal(got_y, want[0])
