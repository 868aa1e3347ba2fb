"""The Python package keeps the reference binding's surface and error behaviour
(bindings/python/tests/test_operations.py, test_convert.py, test_image.py read the same way).
Argument validation runs on the CPU; anything that produces pixels needs the GPU."""
import numpy as np
import pytest

import slimg_b200 as slimg
from conftest import binding_image


def make(w, h):
    return slimg.Image(width=w, height=h, data=binding_image(w, h).tobytes())


def px(img, col, row):
    o = (row * img.width + col) * 4
    return tuple(img.data[o:o + 4])


@pytest.fixture
def sample():
    return make(10, 8)


# ---- validation (no GPU) ----------------------------------------------------------------------------

def test_image_buffer_size_is_checked():
    with pytest.raises(ValueError, match="Buffer size mismatch"):
        slimg.Image(width=10, height=8, data=bytes(10))


def test_crop_argument_validation(sample):
    with pytest.raises(ValueError, match="either region or aspect_ratio"):
        slimg.crop(sample, region=(0, 0, 5, 4), aspect_ratio=(1, 1))
    with pytest.raises(ValueError, match="region or aspect_ratio"):
        slimg.crop(sample)
    with pytest.raises(slimg.SlimgError):
        slimg.crop(sample, region=(8, 0, 5, 4))


def test_extend_argument_validation(sample):
    with pytest.raises(ValueError, match="either aspect_ratio or size"):
        slimg.extend(sample, aspect_ratio=(1, 1), size=(20, 20), fill="transparent")
    with pytest.raises(ValueError, match="aspect_ratio or size"):
        slimg.extend(sample, fill="transparent")
    for bad in ((256, 0, 0), (-1, 0, 0)):
        with pytest.raises(ValueError, match="0-255"):
            slimg.extend(sample, size=(20, 20), fill=bad)
    for bad in ("red", (255, 0)):
        with pytest.raises(ValueError, match="Invalid fill"):
            slimg.extend(sample, size=(20, 20), fill=bad)
    with pytest.raises(slimg.SlimgError):
        slimg.extend(sample, size=(5, 8), fill="transparent")


def test_resize_argument_validation(sample):
    with pytest.raises(ValueError, match="exactly one"):
        slimg.resize(sample)
    with pytest.raises(ValueError, match="exactly one"):
        slimg.resize(sample, width=5, height=4)
    with pytest.raises(slimg.SlimgError.Resize):
        slimg.resize(sample, scale=0.0)


def test_convert_argument_validation(sample):
    for q in (101, -1, 80.5):
        with pytest.raises(ValueError, match="quality"):
            slimg.convert(sample, format="png", quality=q)
    with pytest.raises(ValueError, match="Unknown format"):
        slimg.convert(sample, format="bmp")
    assert slimg.Format._resolve("JPG") is slimg.Format.JPEG
    assert slimg.Format.from_path("a/b.JPEG") is slimg.Format.JPEG and slimg.Format.from_path("x.txt") is None
    assert slimg.Format.from_bytes(b"\x89PNG\r\n\x1a\n....") is slimg.Format.PNG
    assert slimg.Format.JPEG.extension == "jpg" and slimg.Format.WEBP.can_encode


# ---- pixels (GPU) -----------------------------------------------------------------------------------

@pytest.mark.gpu
class TestOperations:
    def test_crop(self, sample):
        c = slimg.crop(sample, region=(2, 1, 5, 4))
        assert (c.width, c.height) == (5, 4)
        assert px(slimg.crop(sample, region=(2, 1, 3, 2)), 0, 0)[:2] == (1, 2)
        sq = slimg.crop(sample, aspect_ratio=(1, 1))
        assert sq.width == sq.height
        wide = slimg.crop(make(100, 100), aspect_ratio=(16, 9))
        assert 1.7 < wide.width / wide.height < 1.8

    def test_extend(self, sample):
        e = slimg.extend(sample, aspect_ratio=(1, 1), fill="transparent")
        assert (e.width, e.height) == (10, 10)
        e = slimg.extend(make(6, 10), aspect_ratio=(1, 1), fill="transparent")
        assert (e.width, e.height) == (10, 10)
        img = make(4, 4)
        assert px(slimg.extend(img, size=(6, 6), fill=(255, 0, 0)), 0, 0) == (255, 0, 0, 255)
        assert px(slimg.extend(img, size=(6, 6), fill="transparent"), 0, 0) == (0, 0, 0, 0)
        assert px(slimg.extend(img, size=(6, 6), fill="transparent"), 1, 1) == (0, 0, 0xFF, 0xFF)
        assert px(slimg.extend(img, size=(6, 6), fill=(128, 64, 32, 200)), 0, 0) == (128, 64, 32, 200)
        same = make(10, 10)
        assert slimg.extend(same, aspect_ratio=(1, 1), fill="transparent").data == same.data

    def test_resize(self, sample):
        r = slimg.resize(sample, width=5)
        assert (r.width, r.height) == (5, 4)
        assert slimg.resize(sample, height=4).height == 4
        r = slimg.resize(sample, exact=(20, 20))
        assert (r.width, r.height) == (20, 20)
        r = slimg.resize(sample, fit=(5, 5))
        assert r.width <= 5 and r.height <= 5
        r = slimg.resize(sample, scale=2.0)
        assert (r.width, r.height) == (20, 16)

    def test_convert(self, sample):
        res = slimg.convert(sample, format="png", quality=80, resize=slimg.Resize.width(5))
        assert res.format is slimg.Format.PNG and res.data[:2] == b"\x89P" and (res.width, res.height) == (5, 4)
        back = slimg.decode(res.data)
        assert (back.width, back.height) == (5, 4)
        assert slimg.convert(sample, format="jpeg").data[:2] == b"\xff\xd8"
        res = slimg.convert(make(100, 100), format="png", crop=slimg.Crop.aspect_ratio(16, 9),
                            extend=slimg.Extend.aspect_ratio(1, 1), fill=(255, 255, 255))
        back = slimg.decode(res.data)
        assert back.width == back.height == 100
        with pytest.raises(slimg.SlimgError.Encode):
            slimg.convert(sample, format="jxl")

    def test_jpeg_frontend(self):
        import oracle
        img = make(40, 24)
        co = slimg.jpeg_frontend(img, quality=70)
        want = oracle.jpeg_frontend(np.frombuffer(img.data, np.uint8).reshape(24, 40, 4),
                                    np.array(co.qt_luma, np.uint16), np.array(co.qt_chroma, np.uint16))
        for got, w in zip(co.planes(), want):
            assert np.array_equal(got, w)
