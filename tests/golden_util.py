"""Loaders for tests/golden/*.npz (made by tests/golden/make_golden.py)."""
import hashlib
import os

import numpy as np

from conftest import IMAGE_KINDS

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def jpeg_cases():
    z = np.load(os.path.join(HERE, "jpeg_golden.npz"))
    for i in range(int(z["n"])):
        kind, w, h, q, digest = z[f"case{i}_meta"]
        img = IMAGE_KINDS[str(kind)](int(w), int(h))
        assert _sha(img) == str(digest), "synthetic input generator drifted from the golden file"
        yield dict(id=f"{kind}-{w}x{h}-q{q}", img=img, ql=z[f"case{i}_ql"], qc=z[f"case{i}_qc"],
                   y=z[f"case{i}_y"], cb=z[f"case{i}_cb"], cr=z[f"case{i}_cr"], quality=int(q))


def resize_cases():
    z = np.load(os.path.join(HERE, "resize_golden.npz"))
    for i in range(int(z["n"])):
        kind, w, h, nw, nh, digest = z[f"case{i}_meta"]
        img = IMAGE_KINDS[str(kind)](int(w), int(h))
        assert _sha(img) == str(digest), "synthetic input generator drifted from the golden file"
        yield dict(id=f"{kind}-{w}x{h}-to-{nw}x{nh}", img=img, nw=int(nw), nh=int(nh), out=z[f"case{i}_out"])
