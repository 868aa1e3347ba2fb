"""K2-vfirst across its plan regimes: bands of 1..6 chunks, 1 / 2 / 3 pass-1 cohorts, units of 3 / 4 / 5 blocks, odd and even
strip widths (vplan.cpp picks them from the ratio).  Every case: max 1 LSB against the oracle, noise-image mean bound, and
the kernel that served it (a plan that is rejected falls through to another kernel and would hide a regression here).
Reference behaviour: crates/slimg-core/src/resize.rs:64-90 (Lanczos3 through image 0.25.9)."""
import numpy as np
import pytest

import oracle
from slimg_b200 import _lowlevel as ll
from conftest import noise_image
from test_gpu_parity import _close, _ll_img, _np

# (source w, h, output w, output h): ratios on x / y
SWEEP = [(6000, 416, int(round(6000 / r)), max(8, int(round(416 / r)))) for r in
         (1.3, 1.5, 2.0, 2.5, 3.0, 3.5, 4.0, 4.5, 5.0, 5.5, 6.0, 6.5, 7.0, 7.5, 8.0, 8.5, 9.0, 10.0, 10.9)]
SWEEP += [(6000, 416, 800, 208), (6000, 416, 3000, 60), (4000, 600, 500, 300), (2400, 800, 1000, 90), (1600, 1200, 200, 700),
          (7680, 264, 1920, 66), (5472, 364, 720, 48), (1024, 768, 400, 300), (804, 804, 100, 100),
          # rows that are not 16-byte aligned: re-pitched on the device, then the same kernels (api.cu: run_pipeline_device)
          (5998, 416, 800, 55), (1366, 768, 683, 384), (4001, 300, 801, 60), (2047, 511, 300, 100)]


@pytest.mark.gpu
@pytest.mark.parametrize("sw,sh,ow,oh", SWEEP, ids=lambda v: str(v))
def test_downscale_ratio_sweep(sw, sh, ow, oh):
    img = noise_image(sw, sh, (sw * 31 + oh) % 997)
    want = oracle.resize(img, oracle.ResizeMode.exact(ow, oh))
    got = _np(ll.resize(_ll_img(img), ll.ResizeMode.EXACT(ow, oh)))
    _close(got, want, "noise", f"sweep {sw}x{sh} -> {ow}x{oh}")
    path = ll.PATH_NAMES[ll.last_resize_path()]
    rx, ry = sw / ow, sh / oh
    if max(rx, ry) <= 10.5 and min(rx, ry) > 3.5:
        assert path == "vfirst", f"{rx:.2f} x {ry:.2f} served by {path}"
    assert path in ("vfirst", "tile", "hfirst"), path      # never the f32 fallback for 16-byte-aligned rows
