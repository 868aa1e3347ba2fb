"""Generates the committed golden vectors under tests/golden/.  Run in the build container, where
/root/reference exists and oracle/_ref/libljt_ref.so (the reference's vendored libjpeg-turbo
2.1.5.1) has been built:   python tests/golden/make_golden.py

jpeg_golden.npz   quantised DCT coefficients produced by the REAL libjpeg-turbo (not by the
                  restatement) for small seeded images -- pins oracle and CUDA kernel alike.
resize_golden.npz outputs of the restated image-0.25.9 Lanczos3 (parity unpinned: regression
                  vectors only, the reference cannot be executed here).
Inputs are regenerated from seeds by tests/conftest.py; their SHA-256 is stored beside the output.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import oracle  # noqa: E402
from conftest import IMAGE_KINDS  # noqa: E402

JPEG_CASES = [("noise", 17, 13, 70), ("noise", 33, 47, 70), ("gradient", 64, 64, 70), ("checker", 48, 17, 25),
              ("alpha_ramp", 31, 16, 95), ("noise", 16, 31, 100), ("constant", 9, 9, 50), ("noise", 1, 1, 70)]
RESIZE_CASES = [("noise", 64, 48, 16, 12), ("gradient", 100, 80, 50, 40), ("noise", 16, 16, 64, 64),
                ("checker", 37, 53, 17, 41), ("alpha_ramp", 60, 10, 7, 33), ("noise", 45, 30, 45, 7)]


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    if oracle.ref_lib() is None:
        raise SystemExit("oracle/_ref/libljt_ref.so missing: run `make -C oracle ref` first")
    out = {}
    for i, (kind, w, h, q) in enumerate(JPEG_CASES):
        img = IMAGE_KINDS[kind](w, h)
        ql, qc = oracle.ref_jpeg_quant_table(q, 0), oracle.ref_jpeg_quant_table(q, 1)
        y, cb, cr = oracle.ref_jpeg_frontend(img, ql, qc)
        out[f"case{i}_meta"] = np.array([kind, str(w), str(h), str(q), sha(img)])
        out[f"case{i}_ql"], out[f"case{i}_qc"] = ql, qc
        out[f"case{i}_y"], out[f"case{i}_cb"], out[f"case{i}_cr"] = y, cb, cr
    out["n"] = np.array(len(JPEG_CASES))
    np.savez_compressed(os.path.join(HERE, "jpeg_golden.npz"), **out)

    out = {}
    for i, (kind, w, h, nw, nh) in enumerate(RESIZE_CASES):
        img = IMAGE_KINDS[kind](w, h)
        out[f"case{i}_meta"] = np.array([kind, str(w), str(h), str(nw), str(nh), sha(img)])
        out[f"case{i}_out"] = oracle.resize_exact(img, nw, nh)
    out["n"] = np.array(len(RESIZE_CASES))
    np.savez_compressed(os.path.join(HERE, "resize_golden.npz"), **out)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
