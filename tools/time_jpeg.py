"""Device-resident timing of the JPEG front end (C5: 24 MP batch, q=70)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from slimg_b200 import _lowlevel as ll
w, h, B = 6000, 4000, 32
g = ll.jpeg_geometry(w, h)
ql, qc = ll.jpeg_quant_table(70, 0, 1), ll.jpeg_quant_table(70, 1, 1)
src = torch.randint(0, 256, (B, h, w, 4), dtype=torch.uint8, device="cuda")
y = torch.empty((B, g.y_HB, g.y_WB, 64), dtype=torch.int16, device="cuda")
cb = torch.empty((B, g.c_HB, g.c_WB, 64), dtype=torch.int16, device="cuda"); cr = torch.empty_like(cb)
st = torch.cuda.current_stream().cuda_stream
def step(): ll.jpeg_frontend_batch_device(src.data_ptr(), B, w, h, ql, qc, y.data_ptr(), cb.data_ptr(), cr.data_ptr(), st)
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 20
e0.record()
for _ in range(n): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
byts = (w * h * 4 + (y[0].numel() + 2 * cb[0].numel()) * 2) * B
print(f"c5 jpeg front end: {ms / B * 1e3:.2f} us/image, {B * w * h / ms / 1e6:.1f} Gpx/s, {byts / ms / 1e6:.0f} GB/s algorithmic")
