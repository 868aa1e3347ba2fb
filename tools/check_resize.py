"""Quick GPU parity probe: a few resize / pipeline shapes through the C ABI against the oracle."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
from slimg_b200 import _lowlevel as ll

def run(w, h, resize, crop=None, extend=None, fill=None, order=ll.ORDER_CONVERT, seed=1):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    image = ll.ImageData(w, h, img.tobytes())
    opts = ll.make_options(resize=resize[0], crop=crop[0] if crop else None, extend=extend[0] if extend else None,
                           fill_color=ll.FillColor.SOLID(*fill) if fill else None)
    t0 = time.time()
    out = ll.pipeline(image, opts, order)
    got = np.frombuffer(out.data, np.uint8).reshape(out.height, out.width, 4)
    f = fill or (255, 255, 255, 255)
    if order == ll.ORDER_CONVERT:
        want = oracle.pipeline(img, crop_mode=crop[1] if crop else None, extend_mode=extend[1] if extend else None,
                               fill=f, resize_mode=resize[1])
    else:
        want = oracle.resize(img, resize[1])
        if crop: want = oracle.crop(want, crop[1])
        if extend: want = oracle.extend(want, extend[1], f)
    if got.shape != want.shape:
        print("SHAPE MISMATCH", got.shape, want.shape); return
    err = np.abs(got.astype(int) - want.astype(int))
    bad = np.argwhere(err > 1)
    print(f"{w}x{h} -> {got.shape[1]}x{got.shape[0]} order={order} max={err.max()} mean={err.mean():.5f} nbad={len(bad)}"
          + (f" first bad {bad[:5].tolist()}" if len(bad) else ""))

R, C, E = ll.ResizeMode, ll.CropMode, ll.ExtendMode
oR, oC, oE = oracle.ResizeMode, oracle.CropMode, oracle.ExtendMode
print("launches before", ll.kernel_launches())
run(640, 480, (R.WIDTH(160), oR.width(160)))
run(600, 400, (R.WIDTH(80), oR.width(80)))
run(4000, 3000, (R.WIDTH(800), oR.width(800)))
run(1000, 336, (R.WIDTH(300), oR.width(300)))
run(604, 400, (R.EXACT(100, 300), oR.exact(100, 300)))
run(6000, 4000, (R.WIDTH(800), oR.width(800)), (C.ASPECT_RATIO(16, 9), oC.aspect_ratio(16, 9)),
    (E.ASPECT_RATIO(1, 1), oE.aspect_ratio(1, 1)), (255, 255, 255, 255), ll.ORDER_CHAIN)
run(6000, 4000, (R.WIDTH(800), oR.width(800)), (C.ASPECT_RATIO(16, 9), oC.aspect_ratio(16, 9)),
    (E.ASPECT_RATIO(1, 1), oE.aspect_ratio(1, 1)), (10, 200, 30, 255), ll.ORDER_CONVERT)
run(7680, 4320, (R.WIDTH(1920), oR.width(1920)))
print("launches after", ll.kernel_launches())
