"""Build recipe of libslimg_b200.so: nvcc, sm_100a only, in-tree (the .so travels to the GPU box)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libslimg_b200.so")
SOURCES = ["geometry.cpp", "taps.cpp", "fastplan.cpp", "vplan.cpp", "tcplan.cpp", "tileplan.cpp", "k1_remap.cu", "k2_resize.cu", "k2_fast.cu", "k2_vfirst.cu", "k2_tc.cu", "k2_tile.cu", "k3_jpeg.cu", "k4_planes.cu", "api.cu", "uniffi_abi.cpp"]
HEADERS = ["host.h", "kernels.cuh", os.path.join("..", "..", "include", "slimg_b200.h"),
           os.path.join("..", "..", "include", "slimg_uniffi.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-fno-fast-math,-Wall",
    "--shared", "-cudart", "static", "-x", "cu",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    extra = os.environ.get("SLIMG_B200_NVCC_EXTRA", "").split()       # experiments: e.g. -DSLIMG_VF_WARPS=16
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB_PATH]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode:
        raise RuntimeError("nvcc failed building libslimg_b200.so")
    build_jpeg_handoff()
    return LIB_PATH


def build_jpeg_handoff() -> None:
    """libslimg_jpeg_handoff.so (slimg_b200/jpeg_handoff/): the hand-off of the GPU's coefficient planes to a libjpeg
    entropy coder.  Needs a libjpeg: JPEG_INC / JPEG_LIB in the environment, or the libjpeg-turbo sources the
    reference vendors (present in the build container; the built .so travels to the GPU box)."""
    mk = os.path.join(HERE, "jpeg_handoff")
    ref = os.environ.get("SLIMG_REFERENCE", "/root/reference")
    if os.environ.get("JPEG_INC"):
        subprocess.run(["make", "-C", mk, "-s"], check=False)
    elif os.path.isdir(os.path.join(ref, "crates", "libjxl-sys", "libjxl", "third_party", "libjpeg-turbo")):
        subprocess.run(["make", "-C", mk, "-s", "-j8", f"REF={ref}"], check=False)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
