import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
from slimg_b200 import _lowlevel as ll
w, h, tw = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(1)
img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
out = ll.pipeline(ll.ImageData(w, h, img.tobytes()), ll.make_options(resize=ll.ResizeMode.WIDTH(tw)), ll.ORDER_CONVERT)
got = np.frombuffer(out.data, np.uint8).reshape(out.height, out.width, 4)
want = oracle.resize(img, oracle.ResizeMode.width(tw))
err = np.abs(got.astype(int) - want.astype(int))
print(w, h, tw, "max", err.max(), "mean", err.mean())
