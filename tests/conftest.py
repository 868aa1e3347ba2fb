import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _have_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


# ---- parity report: every resize comparison of the GPU session (max / mean abs error, kernel that served it) -----------
PARITY_REPORT = []


@pytest.fixture(scope="session", autouse=True)
def _write_parity_report():
    yield
    if not PARITY_REPORT:
        return
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_report.txt"), "w") as fh:
        fh.write("# case | input kind | out WxH | max abs err | mean abs err | kernel path\n")
        for label, kind, w, h, mx, mean, path in PARITY_REPORT:
            fh.write(f"{label} | {kind} | {w}x{h} | {mx} | {mean:.6f} | {path}\n")


# ---- synthetic inputs ------------------------------------------------------------

def gradient_image(w: int, h: int) -> np.ndarray:
    """The reference's Rust fixture: R=x*255/w, G=y*255/h, B=128, A=255 (tests/integration.rs:3-16)."""
    img = np.empty((h, w, 4), np.uint8)
    img[..., 0] = (np.arange(w, dtype=np.int64) * 255 // w).astype(np.uint8)[None, :]
    img[..., 1] = (np.arange(h, dtype=np.int64) * 255 // h).astype(np.uint8)[:, None]
    img[..., 2] = 128
    img[..., 3] = 255
    return img


def binding_image(w: int, h: int) -> np.ndarray:
    """The bindings' fixture: R=row, G=col, B=A=0xFF (bindings/python/tests/conftest.py:5-15)."""
    img = np.empty((h, w, 4), np.uint8)
    img[..., 0] = (np.arange(h) & 0xFF).astype(np.uint8)[:, None]
    img[..., 1] = (np.arange(w) & 0xFF).astype(np.uint8)[None, :]
    img[..., 2] = 0xFF
    img[..., 3] = 0xFF
    return img


def noise_image(w: int, h: int, index: int = 0) -> np.ndarray:
    """Uniform random bytes in all four channels, seed 0x51136 + image index."""
    rng = np.random.default_rng(0x51136 + index)
    return rng.integers(0, 256, (h, w, 4), dtype=np.uint8)


def checker_image(w: int, h: int) -> np.ndarray:
    yy, xx = np.mgrid[0:h, 0:w]
    v = (((yy + xx) & 1) * 255).astype(np.uint8)
    return np.repeat(v[..., None], 4, axis=2).copy()


def constant_image(w: int, h: int, v: int = 128) -> np.ndarray:
    return np.full((h, w, 4), v, np.uint8)


def alpha_ramp_image(w: int, h: int) -> np.ndarray:
    img = noise_image(w, h, 7)
    img[..., 3] = (np.arange(w) * 255 // max(w - 1, 1)).astype(np.uint8)[None, :]
    return img


IMAGE_KINDS = {
    "gradient": gradient_image, "noise": noise_image, "checker": checker_image,
    "constant": constant_image, "alpha_ramp": alpha_ramp_image,
}
