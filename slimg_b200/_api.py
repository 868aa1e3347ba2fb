"""User-facing Python API of slimg_b200.

Same names, argument meaning and error behaviour as the reference's Python package
(bindings/python/slimg/_types.py: Image :137-187, Resize/Crop/Extend :220-274, convert :343-380,
crop_image :383-406, extend_image :409-434, resize_image :437-477), so the reference's binding
tests read the same here (tests/test_binding_api.py).  Pixel work runs on the GPU through
``_lowlevel``; encoding to a file format is outside the hot path (entropy
coding stays on the reference libraries) and is delegated to a pluggable host encoder.
"""
from __future__ import annotations

import builtins
import enum
import io
from typing import Callable, Dict, Optional, Sequence, Tuple, Union

from . import _lowlevel as ll

SlimgError = ll.SlimgError


class Format(enum.Enum):
    JPEG = "jpeg"
    PNG = "png"
    WEBP = "webp"
    AVIF = "avif"
    JXL = "jxl"
    QOI = "qoi"

    @property
    def extension(self) -> str:
        return ll.format_extension(_TO_LL[self])

    @property
    def can_encode(self) -> bool:
        return ll.format_can_encode(_TO_LL[self])

    @classmethod
    def from_path(cls, path: str) -> Optional["Format"]:
        f = ll.format_from_extension(path)
        return None if f is None else _FROM_LL[f]

    @classmethod
    def from_bytes(cls, data: bytes) -> Optional["Format"]:
        f = ll.format_from_magic_bytes(data)
        return None if f is None else _FROM_LL[f]

    @classmethod
    def _resolve(cls, fmt: Union["Format", str]) -> "Format":
        if isinstance(fmt, cls):
            return fmt
        if not isinstance(fmt, str):
            raise TypeError(f"Expected Format or str, got {type(fmt)}")
        key = fmt.strip().lower()
        key = "jpeg" if key == "jpg" else key
        for member in cls:
            if member.value == key:
                return member
        raise ValueError(f"Unknown format: {fmt}")


_TO_LL = {Format.JPEG: ll.Format.JPEG, Format.PNG: ll.Format.PNG, Format.WEBP: ll.Format.WEB_P,
          Format.AVIF: ll.Format.AVIF, Format.JXL: ll.Format.JXL, Format.QOI: ll.Format.QOI}
_FROM_LL = {v: k for k, v in _TO_LL.items()}


class Image:
    """Decoded RGBA8 image: ``width * height * 4`` tightly packed bytes (codec/mod.rs:11-39)."""

    __slots__ = ("_width", "_height", "_data", "_format")

    def __init__(self, width: int, height: int, data: bytes, format: Optional[Format] = None):
        need = width * height * 4
        if len(data) != need:
            raise ValueError(f"Buffer size mismatch: expected {need} bytes ({width}x{height} RGBA), got {len(data)}")
        self._width, self._height, self._data, self._format = width, height, data, format

    width = property(lambda self: self._width)
    height = property(lambda self: self._height)
    data = property(lambda self: self._data)
    format = property(lambda self: self._format)

    def _ll(self) -> ll.ImageData:
        return ll.ImageData(self._width, self._height, self._data)

    @classmethod
    def _wrap(cls, img: ll.ImageData, fmt: Optional[Format] = None) -> "Image":
        return cls(img.width, img.height, img.data, fmt)

    @classmethod
    def _from_raw(cls, width: int, height: int, data: bytes) -> "Image":
        return cls(width, height, data)

    def __repr__(self) -> str:
        return f"Image({self._width}x{self._height})"


class Result:
    """Encoded bytes plus their format; ``width``/``height`` are the post-pipeline dimensions
    (pipeline.rs:97-102)."""

    def __init__(self, data: bytes, format: Format, width: Optional[int] = None, height: Optional[int] = None):
        self._data, self._format, self.width, self.height = data, format, width, height

    data = property(lambda self: self._data)
    format = property(lambda self: self._format)

    def save(self, path: str) -> None:
        with builtins.open(path, "wb") as fh:
            fh.write(self._data)


class Resize:
    width = staticmethod(lambda value: ll.ResizeMode.WIDTH(value=value))
    height = staticmethod(lambda value: ll.ResizeMode.HEIGHT(value=value))
    exact = staticmethod(lambda width, height: ll.ResizeMode.EXACT(width=width, height=height))
    fit = staticmethod(lambda max_width, max_height: ll.ResizeMode.FIT(max_width=max_width, max_height=max_height))
    scale = staticmethod(lambda factor: ll.ResizeMode.SCALE(factor=factor))


class Crop:
    region = staticmethod(lambda x, y, width, height: ll.CropMode.REGION(x=x, y=y, width=width, height=height))
    aspect_ratio = staticmethod(lambda width, height: ll.CropMode.ASPECT_RATIO(width=width, height=height))


class Extend:
    aspect_ratio = staticmethod(lambda width, height: ll.ExtendMode.ASPECT_RATIO(width=width, height=height))
    size = staticmethod(lambda width, height: ll.ExtendMode.SIZE(width=width, height=height))


FillSpec = Union[None, str, Tuple[int, int, int], Tuple[int, int, int, int]]


def _channel(v, name: str) -> int:
    if not isinstance(v, int) or isinstance(v, bool) or v < 0 or v > 255:
        raise ValueError(f"{name} must be an integer in 0-255, got {v!r}")
    return v


def _fill(spec: FillSpec) -> ll.FillColor:
    if spec is None or spec == "transparent":
        return ll.FillColor.TRANSPARENT()
    if isinstance(spec, tuple) and len(spec) in (3, 4):
        rgba = [_channel(v, n) for v, n in zip(spec, "rgba")]
        if len(rgba) == 3:
            rgba.append(255)
        return ll.FillColor.SOLID(r=rgba[0], g=rgba[1], b=rgba[2], a=rgba[3])
    raise ValueError(f"Invalid fill: {spec!r}. Use 'transparent', (r, g, b), or (r, g, b, a).")


def _quality(q) -> int:
    if not isinstance(q, int) or isinstance(q, bool) or not 0 <= q <= 100:
        raise ValueError(f"quality must be an integer in 0-100, got {q!r}")
    return q


def _exactly_one(options: Sequence[Tuple[str, object]], either_msg: str, none_msg: str):
    given = [(k, v) for k, v in options if v is not None]
    if len(given) > 1:
        raise ValueError(either_msg)
    if not given:
        raise ValueError(none_msg)
    return given[0]


# ---- standalone operations -----------------------------------------------------------------------

def crop_image(image: Image, *, region=None, aspect_ratio=None) -> Image:
    kind, val = _exactly_one([("region", region), ("aspect_ratio", aspect_ratio)],
                             "Specify either region or aspect_ratio, not both", "Specify region or aspect_ratio")
    mode = Crop.region(*val) if kind == "region" else Crop.aspect_ratio(*val)
    return Image._wrap(ll.crop(image._ll(), mode), image.format)


def extend_image(image: Image, *, aspect_ratio=None, size=None, fill: FillSpec = "transparent") -> Image:
    kind, val = _exactly_one([("aspect_ratio", aspect_ratio), ("size", size)],
                             "Specify either aspect_ratio or size, not both", "Specify aspect_ratio or size")
    mode = Extend.aspect_ratio(*val) if kind == "aspect_ratio" else Extend.size(*val)
    return Image._wrap(ll.extend(image._ll(), mode, _fill(fill)), image.format)


def resize_image(image: Image, *, width=None, height=None, exact=None, fit=None, scale=None) -> Image:
    msg = "Specify exactly one of: width, height, exact, fit, scale"
    kind, val = _exactly_one([("width", width), ("height", height), ("exact", exact), ("fit", fit), ("scale", scale)],
                             msg, msg)
    mode = {"width": lambda: Resize.width(val), "height": lambda: Resize.height(val),
            "exact": lambda: Resize.exact(*val), "fit": lambda: Resize.fit(*val),
            "scale": lambda: Resize.scale(val)}[kind]()
    return Image._wrap(ll.resize(image._ll(), mode), image.format)


def convert_pixels(image: Image, *, resize=None, crop=None, extend=None, fill: FillSpec = None,
                   order: str = "convert") -> Image:
    """The pixel half of ``convert``: crop -> extend -> resize in ONE fused GPU launch
    (pipeline.rs:71-89).  ``order="chain"`` runs resize -> crop -> extend instead.  As in the
    reference binding, ``fill`` defaults to transparent when ``extend`` is given."""
    fill_color = _fill(fill) if (fill is not None or extend is not None) else None
    opts = ll.make_options(resize, crop, extend, fill_color)
    which = {"convert": ll.ORDER_CONVERT, "chain": ll.ORDER_CHAIN}[order]
    return Image._wrap(ll.pipeline(image._ll(), opts, which), image.format)


def resize_crop_extend(image: Image, *, resize=None, crop=None, extend=None, fill: FillSpec = None) -> Image:
    """BASELINE.json's headline chain: resize, then crop, then extend, fused."""
    return convert_pixels(image, resize=resize, crop=crop, extend=extend, fill=fill, order="chain")


# ---- encoders: NOT part of the GPU path ----------------------------------------------------------

Encoder = Callable[[Image, int], bytes]
_ENCODERS: Dict[Format, Encoder] = {}


def register_encoder(fmt: Format, fn: Encoder) -> None:
    """Plug a host encoder (the reference's MozJPEG / OxiPNG / libwebp ... binding) in for `fmt`."""
    _ENCODERS[fmt] = fn


def _pillow_encoder(pil_name: str, rgb_only: bool) -> Encoder:
    def encode(img: Image, quality: int) -> bytes:
        try:
            from PIL import Image as PILImage
        except ImportError as exc:  # pragma: no cover
            raise SlimgError.Encode("no host encoder registered and Pillow is not installed") from exc
        pil = PILImage.frombytes("RGBA", (img.width, img.height), img.data)
        if rgb_only:
            pil = pil.convert("RGB")      # drops alpha without compositing, like codec/mod.rs:42-48
        buf = io.BytesIO()
        pil.save(buf, pil_name, quality=max(1, quality))
        return buf.getvalue()
    return encode


def _encode(img: Image, fmt: Format, quality: int) -> bytes:
    enc = _ENCODERS.get(fmt)
    if enc is None and fmt in (Format.JPEG, Format.PNG, Format.WEBP):
        enc = _pillow_encoder({Format.JPEG: "JPEG", Format.PNG: "PNG", Format.WEBP: "WEBP"}[fmt], fmt == Format.JPEG)
    if enc is None:
        raise SlimgError.Encode(f"no host encoder registered for {fmt.value}; entropy coding stays on the "
                                "reference libraries (register one with slimg_b200._api.register_encoder)")
    return enc(img, quality)


def convert(image: Image, format: Union[Format, str], quality: int = 80, *, resize=None, crop=None, extend=None,
            fill: FillSpec = None) -> Result:
    _quality(quality)
    fmt = Format._resolve(format)
    pixels = convert_pixels(image, resize=resize, crop=crop, extend=extend, fill=fill)
    return Result(_encode(pixels, fmt, quality), fmt, pixels.width, pixels.height)


def decode(data: bytes) -> Image:
    """Host-side decode (out of scope for the GPU path; Pillow stands in for the reference codecs)."""
    fmt = Format.from_bytes(data)
    if fmt is None:
        raise SlimgError.UnknownFormat("unrecognised magic bytes")
    try:
        from PIL import Image as PILImage
        pil = PILImage.open(io.BytesIO(data)).convert("RGBA")
    except Exception as exc:
        raise SlimgError.Decode(str(exc)) from exc
    return Image(pil.width, pil.height, pil.tobytes(), fmt)


def open(path: str) -> Image:  # noqa: A001 - mirrors slimg.open
    try:
        with builtins.open(path, "rb") as fh:
            data = fh.read()
    except OSError as exc:
        raise SlimgError.Io(str(exc)) from exc
    return decode(data)


def optimize(data: bytes, quality: int = 80) -> Result:
    _quality(quality)
    img = decode(data)
    return Result(_encode(img, img.format, quality), img.format, img.width, img.height)


def optimize_file(path: str, quality: int = 80) -> Result:
    _quality(quality)
    try:
        with builtins.open(path, "rb") as fh:
            data = fh.read()
    except OSError as exc:
        raise SlimgError.Io(str(exc)) from exc
    return optimize(data, quality)


# ---- JPEG pre-entropy front end ---------------------------------------------------------------------

class JpegCoefficients:
    """Quantised DCT coefficient planes ``[HB][WB][64]`` int16 (natural order) for Y, Cb, Cr."""

    def __init__(self, geom, y: bytes, cb: bytes, cr: bytes, qt_luma, qt_chroma):
        self.geometry, self.y, self.cb, self.cr = geom, y, cb, cr
        self.qt_luma, self.qt_chroma = list(qt_luma), list(qt_chroma)

    def planes(self):
        import numpy as np
        g = self.geometry
        return (np.frombuffer(self.y, np.int16).reshape(g.y_HB, g.y_WB, 64),
                np.frombuffer(self.cb, np.int16).reshape(g.c_HB, g.c_WB, 64),
                np.frombuffer(self.cr, np.int16).reshape(g.c_HB, g.c_WB, 64))


def jpeg_quant_tables(quality: int, tables: str = "annex_k"):
    """(luma, chroma) tables for a quality setting; ``tables`` = 'annex_k' (libjpeg-turbo default)
    or 'mozjpeg' (MozJPEG's default table set 3)."""
    _quality(quality)
    base = {"annex_k": ll.QT_ANNEX_K, "mozjpeg": ll.QT_MOZJPEG}[tables]
    return ll.jpeg_quant_table(quality, 0, base), ll.jpeg_quant_table(quality, 1, base)


def jpeg_frontend(image: Image, quality: int = 80, *, tables: str = "annex_k", qt_luma=None, qt_chroma=None
                  ) -> JpegCoefficients:
    """What JpegCodec::encode (codec/jpeg.rs:51-72) computes before entropy coding."""
    if qt_luma is None or qt_chroma is None:
        qt_luma, qt_chroma = jpeg_quant_tables(quality, tables)
    geom, y, cb, cr = ll.jpeg_frontend(image._ll(), qt_luma, qt_chroma)
    return JpegCoefficients(geom, y, cb, cr, qt_luma, qt_chroma)
