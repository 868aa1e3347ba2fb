"""Device-resident timing of one pipeline geometry (CUDA events, L2 flushed by batch size).
usage: time_config.py NAME   with NAME in c1 | c3 | c3conv | c4a | c4b | c2 | r6 | r55 | r3 | r2 | r15 (other ratios of a 24 MP source)"""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from slimg_b200 import _lowlevel as ll

CFG = {
    "c1": (4000, 3000, dict(resize=ll.ResizeMode.WIDTH(800)), ll.ORDER_CONVERT, 64),
    "c2": (4000, 3000, dict(crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1), fill_color=ll.FillColor.TRANSPARENT()), ll.ORDER_CONVERT, 32),
    "c3": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(800), crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1), fill_color=ll.FillColor.SOLID(255, 255, 255, 255)), ll.ORDER_CHAIN, 64),
    "c3conv": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(800), crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1), fill_color=ll.FillColor.SOLID(255, 255, 255, 255)), ll.ORDER_CONVERT, 64),
    "c4a": (7680, 4320, dict(resize=ll.ResizeMode.WIDTH(1920)), ll.ORDER_CONVERT, 32),
    "r6": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(1000)), ll.ORDER_CONVERT, 64),
    "r55": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(1090)), ll.ORDER_CONVERT, 64),
    "r3": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(2000)), ll.ORDER_CONVERT, 32),
    "r2": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(3000)), ll.ORDER_CONVERT, 32),
    "r15": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(4000)), ll.ORDER_CONVERT, 32),
    "c4b": (1024, 1024, dict(resize=ll.ResizeMode.SCALE(4.0)), ll.ORDER_CONVERT, 64),
    "up2": (2048, 2048, dict(resize=ll.ResizeMode.SCALE(2.0)), ll.ORDER_CONVERT, 32),
    "thumb": (4000, 3000, dict(resize=ll.ResizeMode.FIT(400, 400)), ll.ORDER_CONVERT, 64),          # gui/src-tauri/src/commands.rs:129-133
    "thumb24": (6000, 4000, dict(resize=ll.ResizeMode.FIT(400, 400)), ll.ORDER_CONVERT, 64),
    "odd24": (5998, 4000, dict(resize=ll.ResizeMode.WIDTH(800)), ll.ORDER_CONVERT, 32),          # rows not 16-byte aligned: the f32 fallback
    "r748": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(802)), ll.ORDER_CONVERT, 32),
    "r7": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(856)), ll.ORDER_CONVERT, 32),
    "r8": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(752)), ll.ORDER_CONVERT, 32),
    "r86": (6000, 4000, dict(resize=ll.ResizeMode.WIDTH(700)), ll.ORDER_CONVERT, 32),
    "r57": (4000, 3000, dict(resize=ll.ResizeMode.WIDTH(700)), ll.ORDER_CONVERT, 64),
    "r8b": (4000, 3000, dict(resize=ll.ResizeMode.WIDTH(500)), ll.ORDER_CONVERT, 64),
    "odd2": (1366, 768, dict(resize=ll.ResizeMode.WIDTH(683)), ll.ORDER_CONVERT, 256),
    "odd2a": (1368, 768, dict(resize=ll.ResizeMode.WIDTH(684)), ll.ORDER_CONVERT, 256),
    "thumbs": (1024, 768, dict(resize=ll.ResizeMode.FIT(400, 400)), ll.ORDER_CONVERT, 256),
}
name = sys.argv[1]
w, h, kw, order, B = CFG[name]
opts = ll.make_options(**kw)
ow, oh = ll.pipeline_output_dims(w, h, opts, order)
src = torch.randint(0, 256, (B, h, w, 4), dtype=torch.uint8, device="cuda")
dst = torch.empty((B, oh, ow, 4), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def step(): ll.pipeline_batch_device(src.data_ptr(), B, w, h, opts, order, dst.data_ptr(), dst.numel(), st)
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 20
e0.record()
for _ in range(n): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
inb, outb = w * h * 4, ow * oh * 4
print(f"{name:7s} {os.environ.get('AB', ''):28s} path={ll.PATH_NAMES[ll.last_resize_path()]:8s} {ms / B * 1e3:8.2f} us/image  {B * w * h / ms / 1e6:8.1f} Gpx/s  full-in+out {(inb + outb) * B / ms / 1e6:7.0f} GB/s")
