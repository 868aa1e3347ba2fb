"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol
include/slimg_b200.h declares, reports the reference's error categories before touching the
device, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from slimg_b200 import _build, _lowlevel as ll

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "slimg_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(slimg_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_list_agree():
    assert _declared_symbols() == sorted(ll.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_build.build())
    for name in _declared_symbols():
        assert hasattr(lib, name), name
    assert ll.library().slimg_b200_abi_version() == 2


def test_error_categories_before_device_work():
    img = ll.ImageData(10, 8, bytes(10 * 8 * 4))
    with pytest.raises(ll.SlimgError.Crop):
        ll.crop(img, ll.CropMode.REGION(8, 0, 5, 4))
    with pytest.raises(ll.SlimgError.Extend):
        ll.extend(img, ll.ExtendMode.SIZE(5, 8), ll.FillColor.TRANSPARENT())
    with pytest.raises(ll.SlimgError.Resize):
        ll.resize(img, ll.ResizeMode.SCALE(0.0))
    with pytest.raises(ll.SlimgError.Extend):        # extend.rs:98-104 length check
        ll.extend(ll.ImageData(10, 8, bytes(10 * 8 * 4 + 4)), ll.ExtendMode.SIZE(20, 20), ll.FillColor.TRANSPARENT())
    with pytest.raises(ll.SlimgError.Resize):        # resize.rs:67-75 from_raw failure
        ll.resize(ll.ImageData(10, 8, bytes(16)), ll.ResizeMode.WIDTH(5))
    opts = ll.make_options(resize=ll.ResizeMode.WIDTH(5), crop=ll.CropMode.REGION(0, 0, 50, 4))
    with pytest.raises(ll.SlimgError.Crop):          # pipeline.rs:71-74: crop runs (and fails) first
        ll.pipeline(img, opts, ll.ORDER_CONVERT)
    with pytest.raises(ll.SlimgError.Crop):          # chain order: resize ok, then crop of the 5x4 result fails
        ll.pipeline(img, opts, ll.ORDER_CHAIN)


def test_pipeline_output_dims_both_orders():
    o = ll.make_options(resize=ll.ResizeMode.WIDTH(800), crop=ll.CropMode.ASPECT_RATIO(16, 9),
                        extend=ll.ExtendMode.ASPECT_RATIO(1, 1))
    assert ll.pipeline_output_dims(6000, 4000, o, ll.ORDER_CONVERT) == (800, 800)
    assert ll.pipeline_output_dims(6000, 4000, o, ll.ORDER_CHAIN) == (800, 800)
    o = ll.make_options(crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1))
    assert ll.pipeline_output_dims(4000, 3000, o, ll.ORDER_CONVERT) == (4000, 4000)


def test_quant_tables_match_oracle():
    import oracle
    for q in (1, 25, 50, 70, 95, 100):
        for which in (0, 1):
            for base in (0, 1):
                assert ll.jpeg_quant_table(q, which, base) == list(oracle.jpeg_quant_table(q, which, base))


def test_no_cpu_fallback():
    """Without a CUDA device a compute call must fail loudly, not run somewhere else."""
    if ll.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(ll.SlimgError.Image, match="no CUDA device"):
        ll.crop(ll.ImageData(10, 8, bytes(320)), ll.CropMode.REGION(0, 0, 5, 4))
    assert ll.kernel_launches() == 0


def test_source_window_is_what_the_taps_read():
    """slimg_b200_pipeline_source_window: the rows / columns behind the kept output rectangle (host arithmetic)."""
    # C3 chain order: Width(800) -> 800x533, crop 16:9 keeps rows 41..490, Lanczos3 window of those rows (bench.py)
    o = ll.make_options(resize=ll.ResizeMode.WIDTH(800), crop=ll.CropMode.ASPECT_RATIO(16, 9),
                        extend=ll.ExtendMode.ASPECT_RATIO(1, 1))
    x, y, w, h = ll.pipeline_source_window(6000, 4000, o, ll.ORDER_CHAIN)
    assert (x, w) == (0, 6000) and 3400 < h < 3440 and y > 250
    # convert order: crop 16:9 of 6000x4000 is (0,312,6000,3375); the resize reads exactly that rectangle
    assert ll.pipeline_source_window(6000, 4000, o, ll.ORDER_CONVERT) == (0, 312, 6000, 3375)
    # crop only: the crop region itself (crop.rs:45-66)
    o = ll.make_options(crop=ll.CropMode.REGION(10, 20, 30, 40))
    assert ll.pipeline_source_window(100, 100, o, ll.ORDER_CONVERT) == (10, 20, 30, 40)
    # no stage: the whole image
    assert ll.pipeline_source_window(64, 48, ll.make_options(), ll.ORDER_CONVERT) == (0, 0, 64, 48)
