"""Device-resident time of Width(w) for a sweep of output widths of one source size: finds plan cliffs.
usage: ratio_scan.py [src_w src_h [batch]]   -> one line per output width: ratio, us/image, GB/s (source + output bytes), kernel, plan"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from slimg_b200 import _lowlevel as ll

sw, sh = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (6000, 4000)
B = int(sys.argv[3]) if len(sys.argv) > 3 else 16
src = torch.randint(0, 256, (B, sh, sw, 4), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
ratios = [1.3, 1.5, 1.75, 2.0, 2.3, 2.6, 3.0, 3.3, 3.6, 4.0, 4.3, 4.6, 5.0, 5.3, 5.6, 6.0, 6.3, 6.6, 7.0, 7.2, 7.4, 7.5, 7.6, 7.8, 8.0, 8.3, 8.6, 9.0, 9.5, 10.0, 10.5, 11.0, 12.0, 14.0]
for r in ratios:
    ow_req = max(1, int(round(sw / r)))
    opts = ll.make_options(resize=ll.ResizeMode.WIDTH(ow_req))
    ow, oh = ll.pipeline_output_dims(sw, sh, opts, ll.ORDER_CONVERT)
    dst = torch.empty((B, oh, ow, 4), dtype=torch.uint8, device="cuda")
    step = lambda: ll.pipeline_batch_device(src.data_ptr(), B, sw, sh, opts, ll.ORDER_CONVERT, dst.data_ptr(), dst.numel(), st)
    for _ in range(2): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 8
    e0.record()
    for _ in range(n): step()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n / B * 1e3
    gbs = (sw * sh * 4 + ow * oh * 4) / us / 1e3
    print(f"{sw}x{sh} -> {ow}x{oh}  ratio {sw / ow:5.2f}  {us:7.2f} us/image  {gbs:6.0f} GB/s  path={ll.PATH_NAMES[ll.last_resize_path()]}", flush=True)
    del dst
