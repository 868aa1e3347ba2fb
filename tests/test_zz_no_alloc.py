"""slimg_b200_webp_planes_batch_device_static -- the entry point that allocates nothing (include/slimg_b200.h).

No counterpart in the reference: the variant was added for callers that cannot
call cudaMalloc.  Its scratch is a static __device__ array inside libslimg_b200.so (k4_planes.cu); the check is that it
produces exactly the planes and flags of the existing entry point (slimg_b200_color_planes_batch_device, whose scratch
the caller allocates) -- integer work, so the bar is bit-exact -- on two calls in a row, the second of which must not
see the flags the first one left in the shared array.

Named test_zz_* so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest

from slimg_b200 import _lowlevel as ll

STATIC_SCRATCH_BYTES = 262144            # SLIMG_WEBP_STATIC_SCRATCH_BYTES


def _images(rng, n, w, h, alpha_kind):
    img = rng.integers(0, 256, (n, h, w, 4), dtype=np.uint8)
    if alpha_kind == "opaque":
        img[..., 3] = 255
    elif alpha_kind == "mixed":                      # image 0: a few translucent rows; image 1: opaque; the rest: noise alpha
        img[0, :, :, 3] = 255
        img[0, h // 3: h // 3 + 3, ::5, 3] = 17
        if n > 1:
            img[1, :, :, 3] = 255
    return img


def _run(torch, lib, img, static):
    n, h, w, _ = img.shape
    cw, chh = (w + 1) // 2, (h + 1) // 2
    dev = torch.device("cuda:0")
    src = torch.from_numpy(img).to(dev)
    fill = 0xA5                                      # planes start dirty: every byte must be written by the call
    y = torch.full((n, h, w), fill, dtype=torch.uint8, device=dev)
    u = torch.full((n, chh, cw), fill, dtype=torch.uint8, device=dev)
    v = torch.full((n, chh, cw), fill, dtype=torch.uint8, device=dev)
    a = torch.full((n, h, w), fill, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        flags = torch.full((n,), 0x7F7F7F7F, dtype=torch.int32, device=dev)
        ll._check(lib.slimg_b200_webp_planes_batch_device_static(src.data_ptr(), n, w, h, y.data_ptr(), u.data_ptr(), v.data_ptr(),
                                                                  a.data_ptr(), flags.data_ptr(), stream))
        torch.cuda.synchronize()
        any_alpha = flags.cpu().numpy().astype(np.uint32)
    else:
        rows = n * chh
        scratch = torch.zeros(((rows + 3) // 4 * 4 + 4 * n,), dtype=torch.uint8, device=dev)
        ll._check(lib.slimg_b200_color_planes_batch_device(src.data_ptr(), n, w, h, ll.PLANES_WEBP_YUV420, y.data_ptr(), u.data_ptr(),
                                                            v.data_ptr(), a.data_ptr(), scratch.data_ptr(), stream))
        torch.cuda.synchronize()
        any_alpha = scratch[(rows + 3) // 4 * 4:].cpu().numpy().view(np.uint32).copy()
    return y.cpu().numpy(), u.cpu().numpy(), v.cpu().numpy(), a.cpu().numpy(), any_alpha


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    import torch
    lib = ll.library()
    rng = np.random.default_rng(0x5717)
    # call 1: translucent images (sets flags in the shared array); call 2: a different geometry whose images 0 and 1
    # are opaque -- stale flags or stale row bytes from call 1 would change its chroma and its any_alpha
    cases = [_images(rng, 5, 157, 93, "noise"), _images(rng, 3, 64, 201, "mixed"), _images(rng, 4, 90, 60, "opaque")]
    launches0 = ll.kernel_launches()
    for k, img in enumerate(cases):
        got = _run(torch, lib, img, static=True)
        want = _run(torch, lib, img, static=False)
        for name, g, w in zip(("y", "u", "v", "a", "any_alpha"), got, want):
            assert g.shape == w.shape, (k, name)
            assert np.array_equal(g, w), f"call {k}: plane {name} differs from slimg_b200_color_planes_batch_device"
    assert ll.kernel_launches() - launches0 == 3 * len(cases)       # scan + convert (static scratch), convert alone (caller scratch): the CUDA path ran
    assert list(_run(torch, lib, cases[1], static=True)[4]) == [1, 0, 1]
    assert list(_run(torch, lib, cases[2], static=True)[4]) == [0, 0, 0, 0]


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_oracle():
    import torch
    import oracle
    lib = ll.library()
    img = _images(np.random.default_rng(7), 2, 50, 37, "mixed")
    y, u, v, a, flags = _run(torch, lib, img, static=True)
    for i in range(2):
        want = oracle.webp_yuva(img[i])
        assert np.array_equal(y[i], want[0]) and np.array_equal(u[i], want[1]) and np.array_equal(v[i], want[2])
        assert bool(flags[i]) == (want[3] is not None)


@pytest.mark.gpu
def test_static_scratch_limit_is_enforced():
    import torch
    lib = ll.library()
    n, w = 64, 8
    h = 2 * (STATIC_SCRATCH_BYTES // n)              # n * ceil(h/2) alone fills the array: + 4 n flags is over
    t = torch.zeros((1,), dtype=torch.uint8, device="cuda:0")
    rc = lib.slimg_b200_webp_planes_batch_device_static(t.data_ptr(), n, w, h, t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), None, None)
    assert rc != 0
    assert "static scratch" in lib.slimg_b200_last_error().decode()
    # just inside the limit: accepted (tiny images, real buffers)
    h_ok = 2 * ((STATIC_SCRATCH_BYTES - 4 * n) // n)
    img = _images(np.random.default_rng(1), n, w, h_ok, "opaque")
    got = _run(torch, lib, img, static=True)
    want = _run(torch, lib, img, static=False)
    assert all(np.array_equal(g, x) for g, x in zip(got, want))


def test_static_scratch_symbol_is_exported():
    assert hasattr(ll.library(), "slimg_b200_webp_planes_batch_device_static")
