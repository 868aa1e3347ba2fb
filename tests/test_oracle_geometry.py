"""Pins the oracle's (and the C ABI's) region arithmetic on every exact tuple the reference's own
tests hold: crates/slimg-core/src/crop.rs:94-240, extend.rs:141-305, resize.rs:101-139,
tests/integration.rs:59-80, bindings/python/tests/test_operations.py:117-139.  Both
implementations are exercised by the same table: `oracle` (the checker) and `_lowlevel` (the
product's host logic -- no GPU needed for geometry)."""
import pytest

import oracle
from slimg_b200 import _lowlevel as ll


class OracleImpl:
    name = "oracle"
    def crop(self, w, h, m): return oracle.calculate_crop_region(w, h, self._crop(m))
    def extend(self, w, h, m): return oracle.calculate_extend_region(w, h, self._ext(m))
    def dims(self, w, h, m): return oracle.calculate_dimensions(w, h, self._rs(m))
    def out_dims(self, w, h, m): return oracle.resize_output_dims(w, h, self._rs(m))
    err = oracle.OracleError
    @staticmethod
    def _crop(m): return oracle.CropMode.region(*m[1:]) if m[0] == "region" else oracle.CropMode.aspect_ratio(*m[1:])
    @staticmethod
    def _ext(m): return oracle.ExtendMode.aspect_ratio(*m[1:]) if m[0] == "aspect" else oracle.ExtendMode.size(*m[1:])
    @staticmethod
    def _rs(m): return getattr(oracle.ResizeMode, m[0])(*m[1:])


class AbiImpl:
    name = "abi"
    def crop(self, w, h, m): return ll.calculate_crop_region(w, h, self._crop(m))
    def extend(self, w, h, m): return ll.calculate_extend_region(w, h, self._ext(m))
    def dims(self, w, h, m): return ll.calculate_dimensions(w, h, self._rs(m))
    def out_dims(self, w, h, m): return ll.resize_output_dims(w, h, self._rs(m))
    err = ll.SlimgError
    @staticmethod
    def _crop(m): return ll.CropMode.REGION(*m[1:]) if m[0] == "region" else ll.CropMode.ASPECT_RATIO(*m[1:])
    @staticmethod
    def _ext(m): return ll.ExtendMode.ASPECT_RATIO(*m[1:]) if m[0] == "aspect" else ll.ExtendMode.SIZE(*m[1:])
    @staticmethod
    def _rs(m): return getattr(ll.ResizeMode, m[0].upper())(*m[1:])


@pytest.fixture(params=[OracleImpl(), AbiImpl()], ids=lambda i: i.name)
def impl(request):
    return request.param


CROP_OK = [  # (img_w, img_h, mode, expected (x, y, w, h))   crop.rs line
    (200, 100, ("region", 10, 20, 50, 30), (10, 20, 50, 30)),    # :94-108
    (200, 100, ("region", 0, 0, 200, 100), (0, 0, 200, 100)),    # :110-124
    (200, 100, ("aspect", 1, 1), (50, 0, 100, 100)),             # :171-183
    (100, 200, ("aspect", 1, 1), (0, 50, 100, 100)),             # :185-197
    (100, 100, ("aspect", 16, 9), (0, 22, 100, 56)),             # :199-214
    (200, 100, ("aspect", 2, 1), (0, 0, 200, 100)),              # :216-228
    (4000, 3000, ("aspect", 16, 9), (0, 375, 4000, 2250)),       # BASELINE C2
    (6000, 4000, ("aspect", 16, 9), (0, 312, 6000, 3375)),       # BASELINE C3
    (10, 8, ("region", 2, 1, 5, 4), (2, 1, 5, 4)),               # test_operations.py:8-11
]
CROP_ERR = [
    (200, 100, ("region", 150, 0, 100, 50)),    # :126-139
    (200, 100, ("region", 0, 0, 0, 50)),        # :141-154
    (200, 100, ("region", 0, 0, 50, 0)),        # :156-169
    (200, 100, ("aspect", 0, 1)),               # :230-241
    (10, 8, ("region", 8, 0, 5, 4)),            # test_operations.py:29-31
    (200, 100, ("region", 4294967295, 0, 2, 2)),  # u32 wrap in the reference -> Crop error here
]
EXTEND_OK = [  # expected (canvas_w, canvas_h, off_x, off_y)   extend.rs line
    (200, 100, ("aspect", 1, 1), (200, 200, 0, 50)),             # :141-153
    (100, 200, ("aspect", 1, 1), (200, 200, 50, 0)),             # :155-167
    (100, 100, ("aspect", 16, 9), (178, 100, 39, 0)),            # :169-181
    (100, 100, ("aspect", 9, 16), (100, 178, 0, 39)),            # :183-195
    (200, 100, ("aspect", 2, 1), (200, 100, 0, 0)),              # :197-209
    (800, 600, ("size", 1000, 1000), (1000, 1000, 100, 200)),    # :226-238
    (800, 600, ("size", 800, 600), (800, 600, 0, 0)),            # :240-252
    (800, 600, ("size", 1000, 600), (1000, 600, 100, 0)),        # :254-266
    (4000, 2250, ("aspect", 1, 1), (4000, 4000, 0, 875)),        # BASELINE C2
    (6000, 3375, ("aspect", 1, 1), (6000, 6000, 0, 1312)),       # BASELINE C3
    (6, 10, ("aspect", 1, 1), (10, 10, 2, 0)),                   # test_operations.py:46-50
]
EXTEND_ERR = [
    (200, 100, ("aspect", 0, 1)),     # :211-222
    (800, 600, ("size", 500, 500)),   # :268-279
    (800, 600, ("size", 700, 600)),   # :281-292
    (800, 600, ("size", 0, 0)),       # :294-305
    (10, 8, ("size", 5, 8)),          # test_operations.py:76-78
]
RESIZE_OK = [  # (w, h, mode, calculate_dimensions, actual output dims)
    (200, 100, ("width", 100), (100, 50), (100, 50)),      # resize.rs:101-107
    (200, 100, ("height", 50), (100, 50), (100, 50)),      # :109-115
    (400, 200, ("fit", 100, 100), (100, 50), (100, 50)),   # :117-123
    (200, 100, ("scale", 0.5), (100, 50), (100, 50)),      # :125-131
    (200, 100, ("exact", 50, 50), (50, 50), (50, 50)),     # :133-139
    (100, 80, ("width", 50), (50, 40), (50, 40)),          # tests/integration.rs:59-80
    (10, 8, ("width", 5), (5, 4), (5, 4)),                 # test_operations.py:117-120
    (10, 8, ("exact", 20, 20), (20, 20), (20, 20)),        # :126-129
    (10, 8, ("scale", 2.0), (20, 16), (20, 16)),           # :136-139
    (4000, 3000, ("width", 800), (800, 600), (800, 600)),  # BASELINE C1
    (6000, 4000, ("width", 800), (800, 533), (800, 533)),  # BASELINE C3 (chain order)
    (6000, 6000, ("width", 800), (800, 800), (800, 800)),  # BASELINE C3 (convert order)
    (7680, 4320, ("width", 1920), (1920, 1080), (1920, 1080)),   # BASELINE C4a
    (1024, 1024, ("scale", 4.0), (4096, 4096), (4096, 4096)),    # BASELINE C4b
    (1000, 334, ("width", 300), (300, 100), (299, 100)),   # image crate re-derives the dims
    (200, 100, ("width", 200), (200, 100), (200, 100)),    # no-op keeps the image
]
RESIZE_ERR = [
    (200, 100, ("scale", 0.0)), (200, 100, ("scale", -1.0)),   # resize.rs:43-45
    (200, 100, ("width", 0)), (200, 100, ("exact", 0, 5)),     # :54-58
    (1000, 10, ("width", 10)),                                 # height rounds to 0 -> error
]


@pytest.mark.parametrize("w,h,mode,want", CROP_OK)
def test_crop_region(impl, w, h, mode, want):
    assert impl.crop(w, h, mode) == want


@pytest.mark.parametrize("w,h,mode", CROP_ERR)
def test_crop_region_errors(impl, w, h, mode):
    with pytest.raises(impl.err) as e:
        impl.crop(w, h, mode)
    if impl.name == "abi":
        assert isinstance(e.value, ll.SlimgError.Crop)
    else:
        assert e.value.status == oracle.ERR_CROP


@pytest.mark.parametrize("w,h,mode,want", EXTEND_OK)
def test_extend_region(impl, w, h, mode, want):
    assert impl.extend(w, h, mode) == want


@pytest.mark.parametrize("w,h,mode", EXTEND_ERR)
def test_extend_region_errors(impl, w, h, mode):
    with pytest.raises(impl.err) as e:
        impl.extend(w, h, mode)
    if impl.name == "abi":
        assert isinstance(e.value, ll.SlimgError.Extend)
    else:
        assert e.value.status == oracle.ERR_EXTEND


@pytest.mark.parametrize("w,h,mode,want,actual", RESIZE_OK)
def test_resize_dimensions(impl, w, h, mode, want, actual):
    assert impl.dims(w, h, mode) == want
    assert impl.out_dims(w, h, mode) == actual


@pytest.mark.parametrize("w,h,mode", RESIZE_ERR)
def test_resize_dimension_errors(impl, w, h, mode):
    with pytest.raises(impl.err) as e:
        impl.dims(w, h, mode)
    if impl.name == "abi":
        assert isinstance(e.value, ll.SlimgError.Resize)
    else:
        assert e.value.status == oracle.ERR_RESIZE


def test_geometry_agrees_on_a_sweep():
    """oracle and product host logic agree on a dense sweep of odd sizes and ratios."""
    import itertools
    sizes = [(1, 1), (17, 13), (1000, 334), (4001, 2999), (6000, 4000), (333, 1000), (65535, 3)]
    ratios = [(1, 1), (16, 9), (9, 16), (4, 3), (3, 2), (21, 9), (1, 100)]
    for (w, h), (rw, rh) in itertools.product(sizes, ratios):
        for o_mode, l_mode, of, lf in (
            (oracle.CropMode.aspect_ratio(rw, rh), ll.CropMode.ASPECT_RATIO(rw, rh),
             oracle.calculate_crop_region, ll.calculate_crop_region),
            (oracle.ExtendMode.aspect_ratio(rw, rh), ll.ExtendMode.ASPECT_RATIO(rw, rh),
             oracle.calculate_extend_region, ll.calculate_extend_region),
        ):
            try:
                a = of(w, h, o_mode)
            except oracle.OracleError:
                a = None
            try:
                b = lf(w, h, l_mode)
            except ll.SlimgError:
                b = None
            assert a == b, (w, h, rw, rh)
        for tw in (1, 7, 300, 800, 5000):
            for o_mode, l_mode in ((oracle.ResizeMode.width(tw), ll.ResizeMode.WIDTH(tw)),
                                   (oracle.ResizeMode.fit(tw, 2 * tw), ll.ResizeMode.FIT(tw, 2 * tw)),
                                   (oracle.ResizeMode.scale(tw / 1000.0), ll.ResizeMode.SCALE(tw / 1000.0))):
                try:
                    a = oracle.resize_output_dims(w, h, o_mode)
                except oracle.OracleError:
                    a = None
                try:
                    b = ll.resize_output_dims(w, h, l_mode)
                except ll.SlimgError:
                    b = None
                assert a == b, (w, h, tw)
