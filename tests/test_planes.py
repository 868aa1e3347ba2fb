"""Colour-plane pre-stages of the WebP / AVIF encoders.  libwebp and ravif are absent from the
reference tree, so parity is UNPINNED: the oracle restates their published arithmetic (oracle/slimg_oracle.c), these tests
pin the oracle to the colour-science values any correct BT.601 conversion must give, and the GPU to the oracle bit for bit."""
import numpy as np
import pytest

import oracle
from conftest import noise_image, gradient_image
from slimg_b200 import _lowlevel as ll


def _solid(rgb, a=255, w=6, h=4):
    img = np.zeros((h, w, 4), np.uint8)
    img[..., :3] = rgb
    img[..., 3] = a
    return img


# ---- oracle against textbook values (no GPU) -----------------------------------------------------------------------------

@pytest.mark.parametrize("rgb,yuv", [((255, 255, 255), (235, 128, 128)), ((0, 0, 0), (16, 128, 128)), ((255, 0, 0), (82, 90, 240)),
                                     ((0, 255, 0), (145, 54, 34)), ((0, 0, 255), (41, 240, 110)), ((128, 128, 128), (126, 128, 128))])
def test_webp_import_gives_bt601_studio_range(rgb, yuv):
    y, u, v, a = oracle.webp_yuva(_solid(rgb))
    assert a is None and (int(y[0, 0]), int(u[0, 0]), int(v[0, 0])) == yuv
    assert (y == yuv[0]).all() and (u == yuv[1]).all() and (v == yuv[2]).all()


def test_webp_import_shapes_alpha_and_odd_sizes():
    img = noise_image(7, 5, 1)
    y, u, v, a = oracle.webp_yuva(img)
    assert y.shape == (5, 7) and u.shape == v.shape == (3, 4) and a is not None and np.array_equal(a, img[..., 3])
    img[..., 3] = 255
    assert oracle.webp_yuva(img)[3] is None
    # a fully transparent 2x2 block is averaged like an opaque one (libwebp: a == 0 takes the plain path)
    t = _solid((200, 10, 30), a=0)
    o = _solid((200, 10, 30), a=255)
    assert np.array_equal(oracle.webp_yuva(t)[1], oracle.webp_yuva(o)[1])
    # partial alpha: the chroma of a block is pulled towards its opaque pixels
    m = _solid((0, 0, 0), a=255, w=2, h=2)
    m[0, 0] = (255, 0, 0, 255); m[0, 1] = (0, 0, 255, 1); m[1, 0] = (0, 0, 255, 1); m[1, 1] = (0, 0, 255, 1)
    _, u, v, _ = oracle.webp_yuva(m)
    assert int(v[0, 0]) > 200          # red dominates: the three blue pixels are nearly transparent


@pytest.mark.parametrize("rgb,ycc8", [((255, 255, 255), (255, 128, 128)), ((0, 0, 0), (0, 128, 128)), ((255, 0, 0), (76, 85, 255)),
                                      ((0, 0, 255), (29, 255, 107))])
def test_avif_planes_are_bt601_full_range(rgb, ycc8):
    y, cb, cr, a = oracle.avif_planes(_solid(rgb, a=77), 8)
    assert (int(y[0, 0]), int(cb[0, 0]), int(cr[0, 0])) == ycc8 and int(a[0, 0]) == 77
    y, cb, cr, a = oracle.avif_planes(_solid(rgb, a=255), 10)
    assert abs(int(y[0, 0]) - round(ycc8[0] * 1023 / 255)) <= 1 and int(a[0, 0]) == 1023


# ---- GPU against the oracle, bit for bit ---------------------------------------------------------------------------------

def _ll_img(a):
    return ll.ImageData(a.shape[1], a.shape[0], a.tobytes())


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,alpha", [(64, 48, "noise"), (7, 5, "noise"), (641, 479, "opaque"), (1000, 334, "some"), (2, 2, "noise"), (1, 1, "opaque")])
def test_webp_planes_bit_exact(w, h, alpha):
    img = noise_image(w, h, 3)
    if alpha == "opaque":
        img[..., 3] = 255
    elif alpha == "some":
        img[..., 3] = 255
        img[h // 3: h // 3 + 9, w // 4: w // 2, 3] = noise_image(w, h, 4)[h // 3: h // 3 + 9, w // 4: w // 2, 3]
    got = ll.pipeline_planes(_ll_img(img), ll.make_options(), ll.PLANES_WEBP_YUV420)
    y, u, v, a = oracle.webp_yuva(img)
    assert (got.width, got.height, got.chroma_width, got.chroma_height) == (w, h, (w + 1) // 2, (h + 1) // 2)
    assert np.array_equal(np.frombuffer(got.y, np.uint8).reshape(h, w), y)
    assert np.array_equal(np.frombuffer(got.cb, np.uint8).reshape(u.shape), u)
    assert np.array_equal(np.frombuffer(got.cr, np.uint8).reshape(v.shape), v)
    assert (got.a is None) == (a is None)
    if a is not None:
        assert np.array_equal(np.frombuffer(got.a, np.uint8).reshape(h, w), a)


@pytest.mark.gpu
@pytest.mark.parametrize("depth,kind", [(8, ll.PLANES_AVIF_YCBCR444_8), (10, ll.PLANES_AVIF_YCBCR444_10)])
@pytest.mark.parametrize("w,h", [(64, 48), (333, 257), (3, 1)])
def test_avif_planes_bit_exact(depth, kind, w, h):
    img = noise_image(w, h, 5)
    got = ll.pipeline_planes(_ll_img(img), ll.make_options(), kind)
    dt = np.uint8 if depth == 8 else np.uint16
    want = oracle.avif_planes(img, depth)
    for name, g, wv in zip("y cb cr a".split(), (got.y, got.cb, got.cr, got.a), want):
        assert np.array_equal(np.frombuffer(g, dt).reshape(h, w), wv.astype(dt)), name


@pytest.mark.gpu
def test_planes_follow_the_pixel_stages_on_the_device():
    """convert(.., WebP) / convert(.., Avif) up to the encoder: crop -> extend -> resize -> planes, pixels stay in HBM; the
    planes equal the import of the pixels `pipeline` returns.  Thumbnail shape of the GUI: Fit(400, 400)
    (gui/src-tauri/src/commands.rs:129-133)."""
    img = gradient_image(1600, 1200)
    img[..., 2] = noise_image(1600, 1200, 6)[..., 2]
    opts = ll.make_options(resize=ll.ResizeMode.FIT(400, 400), crop=ll.CropMode.ASPECT_RATIO(1, 1))
    px = ll.pipeline(_ll_img(img), opts, ll.ORDER_CONVERT)
    pxa = np.frombuffer(px.data, np.uint8).reshape(px.height, px.width, 4)
    before = ll.kernel_launches()
    got = ll.pipeline_planes(_ll_img(img), opts, ll.PLANES_WEBP_YUV420)
    assert ll.kernel_launches() - before == 2          # K2, planes (the per-block alpha rule needs no scan pass)
    y, u, v, a = oracle.webp_yuva(pxa)
    assert (got.width, got.height) == (400, 400)
    assert np.array_equal(np.frombuffer(got.y, np.uint8).reshape(400, 400), y) and np.array_equal(np.frombuffer(got.cb, np.uint8).reshape(200, 200), u)
    got = ll.pipeline_planes(_ll_img(img), opts, ll.PLANES_AVIF_YCBCR444_10)
    want = oracle.avif_planes(pxa, 10)
    assert np.array_equal(np.frombuffer(got.cr, np.uint16).reshape(400, 400), want[2])
