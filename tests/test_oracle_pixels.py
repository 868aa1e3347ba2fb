"""The oracle's pixel movers and resize against every pixel-level expectation the reference's tests
hold (crop.rs:281-327, extend.rs:333-407, test_operations.py:12-16,52-84) and against an independent
float64 restatement of image-0.25.9's Lanczos3 (resize pixels are otherwise unpinned)."""
import numpy as np
import pytest

import oracle
from conftest import binding_image, constant_image, gradient_image, noise_image
from golden_util import resize_cases
from oracle import CropMode, ExtendMode, ResizeMode


def test_crop_preserves_pixel_data():            # crop.rs:281-309
    img = np.zeros((2, 4, 4), np.uint8)
    for y in range(2):
        for x in range(4):
            img[y, x] = (x, y, 0, 255)
    out = oracle.crop(img, CropMode.region(1, 0, 2, 1))
    assert out.shape == (1, 2, 4)
    assert out.reshape(-1)[0] == 1 and out.reshape(-1)[4] == 2


def test_crop_full_image_returns_clone():        # crop.rs:311-327
    img = constant_image(100, 50)
    assert np.array_equal(oracle.crop(img, CropMode.region(0, 0, 100, 50)), img)


def test_crop_binding_pixels():                  # test_operations.py:12-16
    out = oracle.crop(binding_image(10, 8), CropMode.region(2, 1, 3, 2))
    assert tuple(out[0, 0][:2]) == (1, 2)


def test_extend_fill_and_preserve():             # extend.rs:333-389
    img = constant_image(2, 2)
    red = oracle.extend(img, ExtendMode.size(4, 4), (255, 0, 0, 255))
    assert tuple(red[0, 0]) == (255, 0, 0, 255)
    clear = oracle.extend(img, ExtendMode.size(4, 4), (0, 0, 0, 0))
    assert tuple(clear[0, 0]) == (0, 0, 0, 0)
    two = np.array([[[10, 20, 30, 255], [40, 50, 60, 255]]], np.uint8)
    out = oracle.extend(two, ExtendMode.size(4, 3), (0, 0, 0, 0))
    assert tuple(out[1, 1]) == (10, 20, 30, 255) and tuple(out[1, 2]) == (40, 50, 60, 255)
    assert out.sum() == two.sum()


def test_extend_noop_when_matching():            # extend.rs:391-407
    img = constant_image(200, 100)
    assert np.array_equal(oracle.extend(img, ExtendMode.aspect_ratio(2, 1), (255, 255, 255, 255)), img)


def test_extend_binding_pixels():                # test_operations.py:52-84
    img = binding_image(4, 4)
    out = oracle.extend(img, ExtendMode.size(6, 6), (0, 0, 0, 0))
    assert tuple(out[0, 0]) == (0, 0, 0, 0) and tuple(out[1, 1]) == (0, 0, 255, 255)
    out = oracle.extend(img, ExtendMode.size(6, 6), (128, 64, 32, 200))
    assert tuple(out[0, 0]) == (128, 64, 32, 200)
    # odd padding: the extra pixel goes right / bottom (extend.rs:64-65 floor division)
    out = oracle.extend(img, ExtendMode.size(7, 9), (9, 9, 9, 9))
    assert tuple(out[2, 1]) == tuple(img[0, 0]) and tuple(out[1, 1]) == (9, 9, 9, 9)


def test_pipeline_16_9_then_1_1_is_square():     # SlimgTest.kt:324-341, test_convert.py:55-65
    out = oracle.pipeline(binding_image(100, 100), crop_mode=CropMode.aspect_ratio(16, 9),
                          extend_mode=ExtendMode.aspect_ratio(1, 1), fill=(255, 255, 255, 255))
    assert out.shape == (100, 100, 4)
    assert tuple(out[0, 0]) == (255, 255, 255, 255) and tuple(out[22, 0]) == tuple(binding_image(100, 100)[22, 0])


def _float64_lanczos(img, nw, nh):
    def axis(inp, out_len):
        n = inp.shape[0]
        ratio = np.float32(n) / np.float32(out_len)
        sr = max(ratio, np.float32(1))
        sup = np.float32(3) * sr
        out = np.zeros((out_len,) + inp.shape[1:], np.float64)
        for o in range(out_len):
            c = (np.float32(o) + np.float32(0.5)) * ratio
            l = int(np.clip(np.floor(c - sup), 0, n - 1))
            r = int(np.clip(np.ceil(c + sup), l + 1, n))
            c = c - np.float32(0.5)
            x = ((np.arange(l, r, dtype=np.float32) - c) / sr).astype(np.float64)
            w = np.where(np.abs(x) < 3, np.sinc(x) * np.sinc(x / 3), 0.0)
            out[o] = np.tensordot(w / w.sum(), inp[l:r], axes=(0, 0))
        return out
    v = axis(img.astype(np.float64), nh)
    h = axis(v.transpose(1, 0, 2), nw).transpose(1, 0, 2)
    return np.floor(np.clip(h, 0, 255) + 0.5).astype(np.uint8)


@pytest.mark.parametrize("w,h,nw,nh", [(300, 400, 80, 60), (64, 64, 256, 256), (53, 37, 17, 41),
                                       (80, 100, 50, 40), (40, 30, 40, 9), (5, 3, 1, 1), (1, 1, 7, 5)])
def test_resize_matches_float64_restatement(w, h, nw, nh):
    img = noise_image(w, h, 3)
    got = oracle.resize_exact(img, nw, nh)
    want = _float64_lanczos(img, nw, nh)
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 1 and d.mean() < 1e-3      # f32 vs f64 can flip a rounding, nothing more


def test_resize_constant_image_stays_constant():   # the reference's own resize fixture (resize.rs:96-99)
    for mode in (ResizeMode.width(100), ResizeMode.scale(2.0), ResizeMode.exact(50, 50)):
        out = oracle.resize(constant_image(200, 100), mode)
        assert (out == 128).all()


def test_resize_dims_follow_image_crate():
    assert oracle.resize(gradient_image(1000, 334), ResizeMode.width(300)).shape == (100, 299, 4)
    same = gradient_image(20, 10)
    assert np.array_equal(oracle.resize(same, ResizeMode.width(20)), same)


def test_resize_tap_table_invariants():
    for n_in, n_out in [(6000, 800), (4000, 533), (1024, 4096), (7680, 1920), (17, 13), (13, 17)]:
        left, count, w = oracle.lanczos3_taps(n_in, n_out)
        assert (left >= 0).all() and (count >= 1).all() and (left + count <= n_in).all()
        assert np.allclose(w.sum(1), 1.0, atol=1e-5)
        assert (np.diff(left) >= 0).all()
    assert oracle.lanczos3_taps(6000, 800)[1].max() == 46       # SURVEY appendix A tap counts
    assert oracle.lanczos3_taps(4000, 800)[1].max() == 31
    assert oracle.lanczos3_taps(1024, 4096)[1].max() == 7


@pytest.mark.parametrize("case", list(resize_cases()), ids=lambda c: c["id"])
def test_resize_golden_vectors(case):
    assert np.array_equal(oracle.resize_exact(case["img"], case["nw"], case["nh"]), case["out"])
