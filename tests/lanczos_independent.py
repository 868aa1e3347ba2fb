"""Two derivations of the reference's Lanczos3 resize that share NO code with oracle/slimg_oracle.c.

Both are written from the algorithm statement in SURVEY.md Appendix A (image 0.25.9 `imageops::resize`: vertical pass
into an un-rounded f32 image, horizontal pass, clamp, round half away from zero), not from the C restatement:

* ``resize_f64``  -- the filter as mathematics: per axis a dense f64 weight matrix, result = Wy @ img @ Wx^T.  It pins
  the oracle's ARITHMETIC against the stated algorithm: the oracle's u8 must be round(f64 value) except where the f64
  value lies within 2^-10 of a rounding tie (f32 accumulation error).
* ``resize_f32``  -- the same loops in strict float32 numpy (elementwise IEEE ops, no FMA), taps accumulated in the
  crate's order, weights through the C library's ``sinf``.  Expected to equal the C oracle bit for bit.

Neither makes the parity "pinned": the crate source is absent.  They remove the risk that the
oracle, the golden file and taps.cpp share one transcription error.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np

_libm = ctypes.CDLL("libm.so.6")
_libm.sinf.restype = ctypes.c_float
_libm.sinf.argtypes = [ctypes.c_float]


def _lanczos3_f64(x: float) -> float:
    if abs(x) >= 3.0:
        return 0.0
    if x == 0.0:
        return 1.0
    a, b = math.pi * x, math.pi * x / 3.0
    return (math.sin(a) / a) * (math.sin(b) / b)


def axis_matrix_f64(in_len: int, out_len: int) -> np.ndarray:
    """[out_len, in_len] f64 weights; support 3 * max(ratio, 1), window [floor(c - s), ceil(c + s)) clipped."""
    ratio = in_len / out_len
    sratio = max(ratio, 1.0)
    support = 3.0 * sratio
    m = np.zeros((out_len, in_len))
    for o in range(out_len):
        c = (o + 0.5) * ratio
        left = min(max(math.floor(c - support), 0), in_len - 1)
        right = min(max(math.ceil(c + support), left + 1), in_len)
        c -= 0.5
        ws = np.array([_lanczos3_f64((i - c) / sratio) for i in range(left, right)])
        m[o, left:right] = ws / ws.sum()
    return m


def resize_f64(img: np.ndarray, nw: int, nh: int) -> np.ndarray:
    """-> f64 [nh, nw, 4], NOT clamped or rounded."""
    h, w = img.shape[:2]
    wy, wx = axis_matrix_f64(h, nh), axis_matrix_f64(w, nw)
    x = img.astype(np.float64)
    v = np.einsum("oh,hwc->owc", wy, x)
    return np.einsum("pw,owc->opc", wx, v)


def _sinc_f32(t: np.float32) -> np.float32:
    if t == np.float32(0.0):
        return np.float32(1.0)
    a = np.float32(t * np.float32(math.pi))          # t * PI_f32
    return np.float32(np.float32(_libm.sinf(a)) / a)


def _lanczos3_f32(x: np.float32) -> np.float32:
    if abs(x) < np.float32(3.0):
        return np.float32(_sinc_f32(x) * _sinc_f32(np.float32(x / np.float32(3.0))))
    return np.float32(0.0)


def axis_taps_f32(in_len: int, out_len: int):
    """Per output index: (left, [f32 weights]) exactly as the crate computes them, every operation in f32."""
    f = np.float32
    ratio = f(f(in_len) / f(out_len))
    sratio = f(1.0) if ratio < f(1.0) else ratio
    support = f(f(3.0) * sratio)
    taps = []
    for o in range(out_len):
        c = f(f(f(o) + f(0.5)) * ratio)
        left = int(min(max(int(np.floor(f(c - support))), 0), in_len - 1))
        right = int(min(max(int(np.ceil(f(c + support))), left + 1), in_len))
        c = f(c - f(0.5))
        ws, s = [], f(0.0)
        for i in range(left, right):
            wv = _lanczos3_f32(f(f(f(i) - c) / sratio))
            ws.append(wv)
            s = f(s + wv)
        taps.append((left, [f(wv / s) for wv in ws]))
    return taps


def _sample_axis0_f32(x: np.ndarray, out_len: int) -> np.ndarray:
    """Resample axis 0 of a float32 array, accumulating tap by tap in f32 (t += px * w)."""
    taps = axis_taps_f32(x.shape[0], out_len)
    out = np.zeros((out_len,) + x.shape[1:], np.float32)
    for o, (left, ws) in enumerate(taps):
        acc = np.zeros(x.shape[1:], np.float32)
        for k, wv in enumerate(ws):
            acc = (acc + x[left + k] * wv).astype(np.float32)
        out[o] = acc
    return out


def resize_f32(img: np.ndarray, nw: int, nh: int) -> np.ndarray:
    """-> u8 [nh, nw, 4]."""
    v = _sample_axis0_f32(img.astype(np.float32), nh)                       # vertical pass, un-rounded f32
    hz = _sample_axis0_f32(np.ascontiguousarray(v.transpose(1, 0, 2)), nw).transpose(1, 0, 2)
    t = np.clip(hz, np.float32(0.0), np.float32(255.0))
    return np.floor(t + np.float32(0.5)).astype(np.uint8)                   # round half away (value >= 0)
