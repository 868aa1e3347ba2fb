"""The oracle's JPEG front end against (a) golden coefficients produced by the reference's vendored
libjpeg-turbo 2.1.5.1 (tests/golden/jpeg_golden.npz) and (b), where oracle/_ref was built, that
library itself on fresh inputs.  Integer pipeline: bit-exact or it is wrong."""
import numpy as np
import pytest

import oracle
from conftest import noise_image
from golden_util import jpeg_cases


@pytest.mark.parametrize("case", list(jpeg_cases()), ids=lambda c: c["id"])
def test_frontend_golden(case):
    y, cb, cr = oracle.jpeg_frontend(case["img"], case["ql"], case["qc"])
    assert np.array_equal(y, case["y"]) and np.array_equal(cb, case["cb"]) and np.array_equal(cr, case["cr"])
    assert np.array_equal(oracle.jpeg_quant_table(case["quality"], 0), case["ql"])
    assert np.array_equal(oracle.jpeg_quant_table(case["quality"], 1), case["qc"])


def test_quality_scaling_q70_is_60_percent():        # jcparam.c:130-152
    t = oracle.jpeg_quant_table(70, 0)
    assert t[0] == (16 * 60 + 50) // 100 and t[63] == (99 * 60 + 50) // 100
    assert (oracle.jpeg_quant_table(100, 0) == 1).all()
    assert oracle.jpeg_quant_table(1, 1).max() == 4950


def test_rgb_ycc_known_answers():                    # jccolor.c:221-235 constants
    assert oracle.rgb_ycc(0, 0, 0) == (0, 128, 128)
    assert oracle.rgb_ycc(255, 255, 255) == (255, 128, 128)
    assert oracle.rgb_ycc(255, 0, 0) == (76, 85, 255)
    assert oracle.rgb_ycc(0, 0, 255) == (29, 255, 107)


def test_fdct_dc_of_flat_block():                    # jfdctint.c: output scaled by 8
    out = oracle.fdct_islow(np.full((8, 8), 100))
    assert out[0, 0] == 100 * 64 and np.count_nonzero(out) == 1


def test_geometry_dummy_blocks():                    # jcmaster.c:149-154, jccoefct.c:286-319
    g = oracle.jpeg_geometry(17, 13)
    assert (g.y_wb, g.y_hb, g.y_WB, g.y_HB, g.c_wb, g.c_hb) == (3, 2, 4, 2, 2, 1)
    g = oracle.jpeg_geometry(6000, 4000)
    assert (g.y_WB, g.y_HB, g.c_WB, g.c_HB) == (750, 500, 375, 250)


@pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("w,h", [(16, 16), (17, 13), (1, 1), (9, 9), (33, 47), (100, 80), (250, 250), (31, 16),
                                 (16, 31), (48, 17), (641, 480), (640, 481)])
def test_frontend_matches_real_libjpeg_turbo(w, h):
    for idx, q in enumerate((70, 100, 25)):
        img = noise_image(w, h, idx)
        ql, qc = oracle.jpeg_quant_table(q, 0), oracle.jpeg_quant_table(q, 1)
        assert np.array_equal(ql, oracle.ref_jpeg_quant_table(q, 0))
        for a, b in zip(oracle.jpeg_frontend(img, ql, qc), oracle.ref_jpeg_frontend(img, ql, qc)):
            assert np.array_equal(a, b)


@pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("w,h", [(16, 16), (53, 37), (129, 65)])
def test_coefficient_handoff_writes_the_same_jpeg(w, h):
    """coefficients handed to the entropy coder with jpeg_write_coefficients (jctrans.c:43) give
    the byte stream the full encoder writes -- the front end can be replaced without the file changing."""
    img = noise_image(w, h, 5)
    ql, qc = oracle.ref_jpeg_quant_table(70, 0), oracle.ref_jpeg_quant_table(70, 1)
    y, cb, cr = oracle.jpeg_frontend(img, ql, qc)
    assert oracle.ref_jpeg_write_coefficients(y, cb, cr, w, h, ql, qc) == oracle.ref_jpeg_encode(img, ql, qc)
    prog = oracle.ref_jpeg_write_coefficients(y, cb, cr, w, h, ql, qc, progressive=True)     # codec/jpeg.rs:61
    assert prog[:3] == b"\xff\xd8\xff"
    assert np.array_equal(oracle.ref_jpeg_decode(prog, w, h), oracle.ref_jpeg_decode(oracle.ref_jpeg_encode(img, ql, qc), w, h))


# ---- MozJPEG's overshoot deringing (restated from memory of mozjpeg-sys 2.2.3: parity unpinned) ----------------------------

def test_deringing_leaves_blocks_without_or_only_maximum_samples_alone():
    rng = np.random.default_rng(3)
    b = rng.integers(-128, 127, (8, 8)).astype(np.int32)          # no sample at 127
    assert np.array_equal(oracle.preprocess_deringing(b, 12), b)
    full = np.full((8, 8), 127, np.int32)
    assert np.array_equal(oracle.preprocess_deringing(full, 12), full)


def test_deringing_overshoots_only_the_maximum_runs_and_within_its_cap():
    b = np.full((8, 8), -100, np.int32)
    b[2:6, 2:6] = 127
    out = oracle.preprocess_deringing(b, 10)
    changed = out != b
    assert changed.any() and (b[changed] == 127).all()              # only clipped samples move
    assert out.max() <= 127 + min(31, 2 * 10) and (out[changed] >= 127).all()
    # the cap follows the DC quantiser and the DC headroom of the block
    assert oracle.preprocess_deringing(b, 2).max() <= 127 + 4
    almost = np.full((8, 8), 127, np.int32)
    almost[0, 0] = 126                                              # headroom (127 * 64 - sum) / count = 1 / 63 = 0
    assert oracle.preprocess_deringing(almost, 50).max() == 127


def test_mozjpeg_front_end_equals_libjpeg_where_nothing_clips():
    img = (noise_image(64, 48, 2) // 2 + 32).astype(np.uint8)       # samples in 32..159: no block reaches 255
    ql, qc = oracle.jpeg_quant_table(70, 0, 1), oracle.jpeg_quant_table(70, 1, 1)
    for a, b in zip(oracle.jpeg_frontend_mozjpeg(img, ql, qc), oracle.jpeg_frontend(img, ql, qc)):
        assert np.array_equal(a, b)
