"""Stages the reference's Python binding -- its wrapper package and its pytest suite, UNMODIFIED -- where the GPU box
can see it: oracle/_ref/pybinding/ (git-ignored like every other artefact made from /root/reference; it travels with
the gpurun snapshot next to oracle/_ref/libljt_ref.so).

TEST INFRASTRUCTURE ONLY.  The wrapper is pure Python (bindings/python/slimg/{__init__,_types}.py); it imports the
UniFFI-generated `slimg._lowlevel`, which tests/test_reference_binding.py replaces with slimg_b200._lowlevel.  Nothing
is edited on the way: the staged files are byte-for-byte copies, checked by hash in the test.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/bindings/python"
DST = os.path.join(HERE, "_ref", "pybinding")
FILES = ["slimg/__init__.py", "slimg/_types.py", "slimg/py.typed", "tests/conftest.py", "tests/test_operations.py",
         "tests/test_convert.py", "tests/test_format.py", "tests/test_decode.py", "tests/test_optimize.py"]


def stage() -> bool:
    if not os.path.isdir(SRC):
        return False
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        if not os.path.exists(src):
            continue
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        with open(src, "rb") as fh:
            manifest[rel] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as fh:
        json.dump({"source": SRC, "sha256": manifest}, fh, indent=1, sort_keys=True)
    return True


if __name__ == "__main__":
    print("staged" if stage() else "no /root/reference here: nothing staged")
