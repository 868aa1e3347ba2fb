#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE config 3: batched resize -> crop -> extend of
synthetic 24 MP RGBA8 images (6000x4000 -> Width(800) -> crop 16:9 -> extend 1:1), input Mpx/s.

A step is one pass of the fused pixel pipeline over one batch of --batch images per GPU
(default 128 = the 8-GPU shard of config 3's 1024 images).  Images are independent, so ranks
shard them with no data-path collective ("scaling": "weak").

  value     device-resident throughput (CUDA events on the launching stream, max over ranks)
  e2e       same metric through the host-buffer C ABI (slimg_b200_pipeline_batch_host): pinned
            host inputs, H2D + kernel + D2H inside the timed region
  roofline  algorithmic HBM bytes per launch / measured launch time, against MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (restated reference path) on this box's host cores, bounded sample, at
                --jobs = all cores and --jobs = 1 (cli/src/commands/mod.rs:64-72)
  configs   (N = 1 only) every other BASELINE config -- C1, C2, C3 in the reference's convert order, C4a, C4b, C5 and
            the fused C3 -> JPEG front end -- timed device-resident the same way, each with its algorithmic bytes,
            roofline fraction and a parity check of one image against the oracle outside the timed region

`--impl reference` times the reference's CPU path (the oracle port: the reference is Rust and
cannot be built on this image) on the same config, rank 0 only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SRC_W, SRC_H = 6000, 4000
MPX_PER_IMAGE = SRC_W * SRC_H / 1e6
# arithmetic of the CUDA path: u8 pixels x s8 weight digits -> s32 on the integer tensor pipe (fixed point, exact),
# f32 only for the final recombination; the CPU reference arm computes in f32 like the image crate
DTYPE = "u8*s8->s32 fixed point"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=128, help="images per GPU per step (device-resident)")
    ap.add_argument("--e2e-batch", type=int, default=32, help="images per GPU per end-to-end step")
    ap.add_argument("--order", default="chain", choices=["chain", "convert"],
                    help="chain = resize->crop->extend (BASELINE metric); convert = the reference's convert() order")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-images", type=int, default=0, help="images in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config table (C1, C2, C4a, C4b, C5 ...)")
    ap.add_argument("--no-pageable", action="store_true", help="skip the pageable-host-buffer e2e leg")
    ap.add_argument("--probe-h2d", action="store_true",
                    help="only run the pinned H2D copy probe (same bytes and stream count as the e2e leg, no kernels)")
    return ap.parse_args()


# ---- workload description ----------------------------------------------------------------------------

def source_rows_needed(order: str):
    """Rows of the 6000x4000 source the pipeline must read (minimal source region)."""
    if order == "convert":       # crop 16:9 (0,312,6000,3375) -> extend 1:1 -> Width(800)
        return 3375
    # chain: Width(800) -> 800x533, crop 16:9 keeps rows 41..490, Lanczos3 window of those rows
    in_len, out_len, first, last = SRC_H, 533, 41, 490
    ratio = np.float32(in_len) / np.float32(out_len)
    support = np.float32(3.0) * max(ratio, np.float32(1.0))
    lo = int(np.clip(np.floor((np.float32(first) + np.float32(0.5)) * ratio - support), 0, in_len - 1))
    c = (np.float32(last) + np.float32(0.5)) * ratio
    hi = int(np.clip(np.ceil(c + support), 0, in_len))
    return hi - lo


def algorithmic_bytes_per_image(order: str) -> int:
    return source_rows_needed(order) * SRC_W * 4 + 800 * 800 * 4


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def build_id() -> str:
    """Hash of everything libslimg_b200.so is built from: an ncu capture belongs to one build."""
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "slimg_b200", "csrc")
    for name in sorted(os.listdir(csrc)) + [os.path.join("..", "_build.py"), os.path.join("..", "..", "include", "slimg_b200.h")]:
        with open(os.path.join(csrc, name), "rb") as fh:
            h.update(name.encode()); h.update(fh.read())
    return h.hexdigest()[:16]


def measured_traffic(order: str, batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of the K2 launch from the committed ncu --set full capture
    (profiles/traffic.json: bytes per image), scaled to this launch's batch.  None unless the capture was taken with
    THIS build of the library (traffic.json records the build id next to the capture)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if t.get("build_id") != build_id():
            return None
        return float(t[order]["dram_bytes_per_image"]) * batch
    except Exception:
        return None


# ---- clocks ------------------------------------------------------------------------------------------

class ClockSampler:
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.gpu)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def wait_first_sample(self, timeout=3.0):
        """nvidia-smi needs a moment to start: do not open the timed region before it samples."""
        t0 = time.time()
        while self.proc is not None and time.time() - t0 < timeout:
            try:
                if os.path.getsize(self.path) > 0:
                    return
            except OSError:
                return
            time.sleep(0.01)

    def stop(self, t_begin=None, t_end=None):
        """Median SM clock and the throttle reasons seen between wall-clock times t_begin and t_end (the timed region,
        widened to the sampling period when it is shorter than that)."""
        import datetime
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        rows = []
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                    rows.append((ts, float(f[1]), float(f[2]), [v.lower().startswith("active") for v in f[5:9]]))
                except ValueError:
                    continue
            os.unlink(self.path)
        except Exception:
            pass
        if t_begin is not None and rows:
            inside = [r for r in rows if t_begin - 0.03 <= r[0] <= t_end + 0.03]
            if not inside:      # nearest sample to the middle of the region
                mid = 0.5 * (t_begin + t_end)
                inside = [min(rows, key=lambda r: abs(r[0] - mid))]
            rows = inside
        if rows:
            out["sm_mhz"] = float(np.median([r[1] for r in rows])); out["sm_max_mhz"] = rows[-1][2]; out["samples"] = len(rows)
            names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
            out["reasons"] = sorted({n for r in rows for n, on in zip(names, r[3]) if on})
        return out


# ---- CPU reference arm -----------------------------------------------------------------------------------

def cpu_pipeline_one(img, order):
    import oracle
    R, Cm, E = oracle.ResizeMode, oracle.CropMode, oracle.ExtendMode
    if order == "convert":
        return oracle.pipeline(img, crop_mode=Cm.aspect_ratio(16, 9), extend_mode=E.aspect_ratio(1, 1),
                               fill=(255, 255, 255, 255), resize_mode=R.width(800))
    out = oracle.resize(img, R.width(800))
    out = oracle.crop(out, Cm.aspect_ratio(16, 9))
    return oracle.extend(out, E.aspect_ratio(1, 1), (255, 255, 255, 255))


def cpu_run(order: str, n_images: int, threads: int, distinct: int = 4):
    """Times the oracle (the restated reference CPU path, one image per thread like rayon's
    par_iter over files) on n_images synthetic 24 MP images.  Returns (seconds, Mpx/s)."""
    from concurrent.futures import ThreadPoolExecutor
    rng = np.random.default_rng(0x51136)
    imgs = [rng.integers(0, 256, (SRC_H, SRC_W, 4), dtype=np.uint8) for _ in range(distinct)]
    cpu_pipeline_one(imgs[0][:400, :600].copy(), "chain")     # load the library
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda i: cpu_pipeline_one(imgs[i % distinct], order).shape, range(n_images)))
    dt = time.perf_counter() - t0
    return dt, n_images * MPX_PER_IMAGE / dt


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def run_reference(args, rank: int):
    if rank != 0:
        return
    threads = host_threads()
    per_step = max(threads, 8)
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_run(args.order, min(per_step, threads), threads)
    total_t, total_n = 0.0, 0
    for _ in range(args.steps):
        dt, _ = cpu_run(args.order, per_step, threads)
        total_t += dt; total_n += per_step
        if total_t > 150:
            break
    value = total_n * MPX_PER_IMAGE / total_t
    sample = f"{per_step} synthetic 24 MP RGBA8 images per step, {threads} threads, one image per thread"
    _emit(json.dumps({
        "impl": "reference", "metric": "Mpx/s resize->crop->extend", "value": value, "unit": "Mpx/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total_t / max(1, total_n // per_step), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, per_step),
        "cpu_baseline": {"value": value, "unit": "Mpx/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mpx/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Rust (no cargo on this image): timed the C restatement of its CPU path (oracle/)",
    }))


def workload_config(args, batch):
    return {"workload": f"C3 batched resize->crop->extend, 6000x4000 RGBA8 -> Width(800) -> crop 16:9 -> extend 1:1 "
                        f"(order={args.order})", "images_per_gpu_per_step": batch, "order": args.order,
            "l2_policy": "inputs larger than L2 (batch >> 126 MB), no flush needed", "parallelism": f"dp{args.gpus}"}



# ---- the other BASELINE configs (N = 1): device-resident, CUDA events, one oracle check each ---------------------------

def config_table(ll, torch, peak, main_src):
    """C1, C2, C3 in convert order, C4a, C4b, C5 (+ un-quantised output) and C3 fused with the JPEG front end: time per
    image over 10 back-to-back launches after 3 warm-ups (inputs are far larger than L2), algorithmic bytes as
    SURVEY.md 8(d) defines them, and parity of the batch's first image against the oracle (outside the timed region)."""
    import oracle
    R, Cm, E = oracle.ResizeMode, oracle.CropMode, oracle.ExtendMode
    WHITE = (255, 255, 255, 255)
    stream = torch.cuda.current_stream().cuda_stream
    out = {}

    def timed(fn, n=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    def err(got, want):
        d = np.abs(got.astype(np.int16) - want.astype(np.int16))
        return {"max_abs_err": int(d.max()), "mean_abs_err": float(d.mean())}

    def pixel_config(name, w, h, kw, order, batch, alg_bytes, oracle_fn, src=None):
        opts = ll.make_options(**kw)
        ow, oh = ll.pipeline_output_dims(w, h, opts, order)
        if src is None:
            gen = torch.Generator(device="cuda").manual_seed(0x51136 + len(out))
            src = torch.randint(0, 256, (batch, h, w, 4), dtype=torch.uint8, device="cuda", generator=gen)
        dst = torch.empty((batch, oh, ow, 4), dtype=torch.uint8, device="cuda")
        ms = timed(lambda: ll.pipeline_batch_device(src.data_ptr(), batch, w, h, opts, order, dst.data_ptr(), dst.numel(), stream))
        path = ll.PATH_NAMES[ll.last_resize_path()]
        us = ms / batch * 1e3
        rec = {"us_per_image": us, "mpx_per_s": w * h / us, "algorithmic_bytes": alg_bytes,
               "frac": alg_bytes / (us * 1e-6) / 1e9 / peak, "batch": batch, "kernel": path}
        rec.update(err(dst[0].cpu().numpy(), oracle_fn(src[0].cpu().numpy())))
        out[name] = rec
        del dst
        return src

    pixel_config("C1 4000x3000 -> Width(800)", 4000, 3000, dict(resize=ll.ResizeMode.WIDTH(800)), ll.ORDER_CONVERT, 64,
                 48_000_000 + 1_920_000, lambda a: oracle.resize(a, R.width(800)))
    pixel_config("C2 4000x3000 crop 16:9 + extend 1:1 transparent", 4000, 3000,
                 dict(crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1), fill_color=ll.FillColor.TRANSPARENT()),
                 ll.ORDER_CONVERT, 32, 36_000_000 + 64_000_000,
                 lambda a: oracle.pipeline(a, crop_mode=Cm.aspect_ratio(16, 9), extend_mode=E.aspect_ratio(1, 1), fill=(0, 0, 0, 0)))
    c3 = dict(resize=ll.ResizeMode.WIDTH(800), crop=ll.CropMode.ASPECT_RATIO(16, 9), extend=ll.ExtendMode.ASPECT_RATIO(1, 1),
              fill_color=ll.FillColor.SOLID(*WHITE))
    pixel_config("C3 convert order (crop 16:9 -> extend 1:1 -> Width(800))", SRC_W, SRC_H, c3, ll.ORDER_CONVERT, 64,
                 algorithmic_bytes_per_image("convert"),
                 lambda a: oracle.pipeline(a, crop_mode=Cm.aspect_ratio(16, 9), extend_mode=E.aspect_ratio(1, 1), fill=WHITE, resize_mode=R.width(800)),
                 src=main_src[:64])
    pixel_config("C4a 7680x4320 -> Width(1920)", 7680, 4320, dict(resize=ll.ResizeMode.WIDTH(1920)), ll.ORDER_CONVERT, 32,
                 132_710_400 + 8_294_400, lambda a: oracle.resize(a, R.width(1920)))
    pixel_config("C4b 1024x1024 -> Scale(4.0)", 1024, 1024, dict(resize=ll.ResizeMode.SCALE(4.0)), ll.ORDER_CONVERT, 64,
                 4_194_304 + 67_108_864, lambda a: oracle.resize(a, R.scale(4.0)))

    pixel_config("GUI thumbnail 4000x3000 -> Fit(400,400)", 4000, 3000, dict(resize=ll.ResizeMode.FIT(400, 400)), ll.ORDER_CONVERT, 64,
                 48_000_000 + 400 * 300 * 4, lambda a: oracle.resize(a, R.fit(400, 400)))

    # rows that are not 16-byte aligned (width % 4 != 0): one K1 copy into a padded pitch, then the same tensor pipeline
    pixel_config("5998x4000 -> Width(800) (rows not 16-byte aligned: re-pitched on the device)", 5998, 4000, dict(resize=ll.ResizeMode.WIDTH(800)),
                 ll.ORDER_CONVERT, 32, 5998 * 4000 * 4 + 800 * 534 * 4, lambda a: oracle.resize(a, R.width(800)))

    # colour planes for the WebP / AVIF encoders, 24 MP batch; planes vs the oracle's restatement (unpinned)
    nb = 16
    srcp = main_src[:nb]
    n_px = SRC_W * SRC_H
    for name, kind, bps, csz in (("WebP import: RGBA -> YUV 4:2:0 + A (libwebp gamma-domain chroma), 6000x4000", ll.PLANES_WEBP_YUV420, 1, n_px // 4),
                                 ("AVIF import: RGBA -> YCbCr 4:4:4 10-bit + A (ravif, BT.601 full range), 6000x4000", ll.PLANES_AVIF_YCBCR444_10, 2, n_px)):
        dt = torch.uint8 if bps == 1 else torch.int16
        py, pa = torch.empty((nb, n_px), dtype=dt, device="cuda"), torch.empty((nb, n_px), dtype=dt, device="cuda")
        pcb, pcr = torch.empty((nb, csz), dtype=dt, device="cuda"), torch.empty((nb, csz), dtype=dt, device="cuda")
        scr = torch.zeros((nb * (SRC_H // 2) + 4 * nb + 64,), dtype=torch.uint8, device="cuda")
        lib = ll.library()
        ms = timed(lambda: ll._check(lib.slimg_b200_color_planes_batch_device(srcp.data_ptr(), nb, SRC_W, SRC_H, kind, py.data_ptr(), pcb.data_ptr(),
                                                                              pcr.data_ptr(), pa.data_ptr(), scr.data_ptr(), stream)))
        img0 = srcp[0].cpu().numpy()
        if kind == ll.PLANES_WEBP_YUV420:
            wy, wu, wv, wa = oracle.webp_yuva(img0)
            diff = int(np.count_nonzero(py[0].cpu().numpy().reshape(wy.shape) != wy)) + int(np.count_nonzero(pcb[0].cpu().numpy().reshape(wu.shape) != wu)) \
                + int(np.count_nonzero(pcr[0].cpu().numpy().reshape(wv.shape) != wv)) + int(np.count_nonzero(pa[0].cpu().numpy().reshape(wy.shape) != wa))
            alg = n_px * 4 + n_px * 2 + 2 * csz
        else:
            want = oracle.avif_planes(img0, 10)
            diff = sum(int(np.count_nonzero(t[0].cpu().numpy().view(np.uint16).reshape(wv.shape) != wv)) for t, wv in zip((py, pcb, pcr, pa), want))
            alg = n_px * 4 + n_px * 2 * 4
        us = ms / nb * 1e3
        out[name] = {"us_per_image": us, "mpx_per_s": n_px / us, "algorithmic_bytes": alg, "frac": alg / (us * 1e-6) / 1e9 / peak, "batch": nb,
                     "kernel": "webp_planes_kernel" if bps == 1 else "avif_planes_kernel", "differing_samples": diff, "parity": "unpinned (libwebp / ravif absent)"}
        del py, pa, pcb, pcr, scr

    # C5: JPEG front end, q = 70, 24 MP batch; MozJPEG's table set (the reference encoder) -- coefficients vs the oracle
    nb = 32
    g = ll.jpeg_geometry(SRC_W, SRC_H)
    ql, qc = ll.jpeg_quant_table(70, 0, ll.QT_MOZJPEG), ll.jpeg_quant_table(70, 1, ll.QT_MOZJPEG)
    mk = lambda hb, wb: torch.empty((nb, hb, wb, 64), dtype=torch.int16, device="cuda")
    y, cb, cr = mk(g.y_HB, g.y_WB), mk(g.c_HB, g.c_WB), mk(g.c_HB, g.c_WB)
    src5 = main_src[:nb]
    ms = timed(lambda: ll.jpeg_frontend_batch_device(src5.data_ptr(), nb, SRC_W, SRC_H, ql, qc, y.data_ptr(), cb.data_ptr(), cr.data_ptr(), stream))
    want = oracle.jpeg_frontend(src5[0].cpu().numpy(), np.array(ql, np.uint16), np.array(qc, np.uint16))
    diff = sum(int(np.count_nonzero(t[0].cpu().numpy() != wv)) for t, wv in zip((y, cb, cr), want))
    us = ms / nb * 1e3
    out["C5 JPEG front end q=70 (RGB->YCbCr 4:2:0, FDCT, quantise), 6000x4000"] = {
        "us_per_image": us, "mpx_per_s": SRC_W * SRC_H / us, "algorithmic_bytes": 96_000_000 + 72_000_000,
        "frac": 168e6 / (us * 1e-6) / 1e9 / peak, "batch": nb, "kernel": "jpeg_frontend_warp_kernel", "differing_coefficients": diff}
    yu, cbu, cru = mk(g.y_HB, g.y_WB), mk(g.c_HB, g.c_WB), mk(g.c_HB, g.c_WB)
    ms = timed(lambda: ll.jpeg_frontend_batch_device_ex(src5.data_ptr(), nb, SRC_W, SRC_H, ql, qc, y.data_ptr(), cb.data_ptr(), cr.data_ptr(),
                                                        yu.data_ptr(), cbu.data_ptr(), cru.data_ptr(), stream))
    wantu = oracle.jpeg_frontend_unquantised(src5[0].cpu().numpy())
    diff = sum(int(np.count_nonzero(t[0].cpu().numpy() != wv)) for t, wv in zip((y, cb, cr, yu, cbu, cru), want + wantu))
    us = ms / nb * 1e3
    out["C5 + un-quantised side output (MozJPEG trellis input)"] = {
        "us_per_image": us, "mpx_per_s": SRC_W * SRC_H / us, "algorithmic_bytes": 96_000_000 + 144_000_000,
        "frac": 240e6 / (us * 1e-6) / 1e9 / peak, "batch": nb, "kernel": "jpeg_frontend_warp_kernel", "differing_coefficients": diff}
    del y, cb, cr, yu, cbu, cru

    # convert(.., Jpeg) on the device: C3 (chain order) and the front end of its 800x800 result, pixels never leave HBM
    nb = 64
    opts = ll.make_options(**c3)
    g8 = ll.jpeg_geometry(800, 800)
    px = torch.empty((nb, 800, 800, 4), dtype=torch.uint8, device="cuda")
    mk = lambda hb, wb: torch.empty((nb, hb, wb, 64), dtype=torch.int16, device="cuda")
    y, cb, cr = mk(g8.y_HB, g8.y_WB), mk(g8.c_HB, g8.c_WB), mk(g8.c_HB, g8.c_WB)
    srcj = main_src[:nb]
    ms = timed(lambda: ll.pipeline_jpeg_batch_device(srcj.data_ptr(), nb, SRC_W, SRC_H, opts, ll.ORDER_CHAIN, ql, qc, px.data_ptr(), px.numel(),
                                                     y.data_ptr(), cb.data_ptr(), cr.data_ptr(), stream=stream))
    want = oracle.jpeg_frontend(px[0].cpu().numpy(), np.array(ql, np.uint16), np.array(qc, np.uint16))
    diff = sum(int(np.count_nonzero(t[0].cpu().numpy() != wv)) for t, wv in zip((y, cb, cr), want))
    us = ms / nb * 1e3
    alg = algorithmic_bytes_per_image("chain") - 800 * 800 * 4 + (g8.y_HB * g8.y_WB + 2 * g8.c_HB * g8.c_WB) * 128
    out["C3 chain -> JPEG front end (convert to JPEG up to the entropy coder, 2 launches)"] = {
        "us_per_image": us, "mpx_per_s": SRC_W * SRC_H / us, "algorithmic_bytes": alg, "frac": alg / (us * 1e-6) / 1e9 / peak,
        "batch": nb, "kernel": "K2 + K3", "differing_coefficients": diff}
    return out


def h2d_ceiling(torch, pin_ptr, image_stride, window_bytes, nb, n_streams, reps, sync_all):
    """The e2e leg's copy pattern with no kernels: nb pinned source windows per step over n_streams streams.
    Returns seconds for `reps` steps (after one warm-up step)."""
    import ctypes as C
    host = np.ctypeslib.as_array(C.cast(pin_ptr, C.POINTER(C.c_uint8)), shape=(nb, image_stride))
    th = torch.from_numpy(host)
    streams = [torch.cuda.Stream() for _ in range(n_streams)]
    dev = [torch.empty(window_bytes, dtype=torch.uint8, device="cuda") for _ in range(n_streams)]

    def one_step():
        for i in range(nb):
            with torch.cuda.stream(streams[i % n_streams]):
                dev[i % n_streams].copy_(th[i, :window_bytes], non_blocking=True)
        torch.cuda.synchronize()

    one_step()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(reps):
        one_step()
    return time.perf_counter() - t0


# ---- B200 arm ------------------------------------------------------------------------------------------

_REAL_STDOUT = None


def _claim_stdout():
    """stdout carries exactly ONE JSON line: everything else a library prints there (NCCL's version banner, torchrun
    notices) is sent to stderr for the length of the run."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _emit(line: str):
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (line + "\n").encode())


def main():
    args = parse_args()
    _claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import ctypes as C
    import torch
    import torch.distributed as dist
    from slimg_b200 import _lowlevel as ll, _shard

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    order = ll.ORDER_CHAIN if args.order == "chain" else ll.ORDER_CONVERT
    opts = ll.make_options(resize=ll.ResizeMode.WIDTH(800), crop=ll.CropMode.ASPECT_RATIO(16, 9),
                           extend=ll.ExtendMode.ASPECT_RATIO(1, 1), fill_color=ll.FillColor.SOLID(255, 255, 255, 255))
    ow, oh = ll.pipeline_output_dims(SRC_W, SRC_H, opts, order)
    B = args.batch

    # synthetic, device-resident batch: uniform random bytes in all four channels
    gen = torch.Generator(device="cuda").manual_seed(0x51136 + rank)
    src = torch.randint(0, 256, (B, SRC_H, SRC_W, 4), dtype=torch.uint8, device="cuda", generator=gen)
    dst = torch.empty((B, oh, ow, 4), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        ll.pipeline_batch_device(src.data_ptr(), B, SRC_W, SRC_H, opts, order, dst.data_ptr(), dst.numel(), stream)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    if rank == 0:
        sampler.wait_first_sample()
    launches0 = ll.kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_begin = time.time()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    t_end = time.time()
    ms = ev0.elapsed_time(ev1)
    launches = ll.kernel_launches() - launches0
    ms_max = _shard.max_over_ranks(ms, device="cuda")
    clocks = sampler.stop(t_begin, t_end) if rank == 0 else None
    value = _shard.whole_job_throughput(B * args.steps * MPX_PER_IMAGE, world, ms_max * 1e-3)

    if args.probe_h2d:
        args.no_cpu = True
        args.no_configs = True
    # roofline of the (single) kernel of the step, from this rank's own launches
    peak, peak_kind = measured_peak_gbs()
    alg_bytes = algorithmic_bytes_per_image(args.order) * B
    launch_ms = ms / max(1, launches)
    achieved = alg_bytes / (launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": measured_traffic(args.order, B), "peak_kind": peak_kind, "kernel": "resize pipeline (K2), path " + ll.PATH_NAMES[ll.last_resize_path()],
                "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": launch_ms}

    # end to end through the host-buffer C ABI
    e2e = None
    if not args.no_e2e:
        lib = ll.library()
        nb = args.e2e_batch
        in_bytes, out_bytes = SRC_W * SRC_H * 4, ow * oh * 4
        # the library uploads only the source window the pipeline reads (slimg_b200_pipeline_source_window)
        wx, wy, ww, wh = ll.pipeline_source_window(SRC_W, SRC_H, opts, order)
        h2d_bytes = ww * wh * 4 if ww * 10 < SRC_W * 9 else wh * SRC_W * 4
        pin_in = lib.slimg_b200_host_alloc(nb * in_bytes)
        pin_out = lib.slimg_b200_host_alloc(nb * out_bytes)
        if not pin_in or not pin_out:
            raise SystemExit("pinned allocation failed")
        host = np.ctypeslib.as_array(C.cast(pin_in, C.POINTER(C.c_uint8)), shape=(nb, in_bytes))
        rng = np.random.default_rng(0x51136 + 1000 + rank)
        block = rng.integers(0, 256, in_bytes, dtype=np.uint8)
        for i in range(nb):
            host[i] = np.roll(block, 4 * 977 * i)
        cin = (ll.CImage * nb)()
        cout = (ll.CImage * nb)()
        st = (C.c_int * nb)()
        dev = (C.c_int32 * 1)(local_rank)
        cfg = ll.CBatchConfig(1, 4, dev)

        def e2e_step():
            for i in range(nb):
                cin[i] = ll.CImage(SRC_W, SRC_H, pin_in + i * in_bytes, in_bytes)
                cout[i] = ll.CImage(0, 0, pin_out + i * out_bytes, out_bytes)
            failed = lib.slimg_b200_pipeline_batch_host(cin, nb, C.byref(opts), order, C.byref(cfg), cout, st)
            if failed:
                raise SystemExit("e2e batch failed: " + lib.slimg_b200_last_error().decode())

        e2e_steps = max(3, min(args.steps, 8))

        def timed_e2e(step_fn, steps):
            for _ in range(3):                # warm-up: stream-ordered pool growth, plan upload, page pinning
                step_fn()
            barrier()
            step_ms = []
            t0 = time.perf_counter()
            for _ in range(steps):
                t1 = time.perf_counter()
                step_fn()
                step_ms.append(round(1e3 * (time.perf_counter() - t1), 2))
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            return _shard.max_over_ranks(dt, device="cuda"), step_ms

        dt_max, step_ms = timed_e2e(e2e_step, e2e_steps)
        e2e_value = _shard.whole_job_throughput(nb * e2e_steps * MPX_PER_IMAGE, world, dt_max)
        checksum = int(np.ctypeslib.as_array(C.cast(pin_out, C.POINTER(C.c_uint8)), shape=(nb * out_bytes,))[::4097].sum())
        e2e = {"value": e2e_value, "unit": "Mpx/s", "h2d_bytes_per_step": nb * h2d_bytes, "d2h_bytes_per_step": nb * out_bytes,
               "host_image_bytes_per_step": nb * in_bytes,
               "images_per_gpu_per_step": nb, "steps": e2e_steps, "timer": "host wall clock around the C-ABI calls, max over ranks",
               "result_checksum": checksum, "step_ms": step_ms, "host_buffers": "pinned (slimg_b200_host_alloc)"}

        # the ceiling of that leg: the same pinned H2D pattern (bytes, 4 streams, all ranks at once) with no kernels
        barrier()
        t_copy = h2d_ceiling(torch, pin_in, in_bytes, h2d_bytes, nb, 4, 3, barrier)
        t_copy = _shard.max_over_ranks(t_copy, device="cuda")
        ceiling = _shard.whole_job_throughput(nb * 3 * MPX_PER_IMAGE, world, t_copy)
        e2e["h2d_copy_ceiling"] = {"value": ceiling, "unit": "Mpx/s", "aggregate_h2d_gbs": world * nb * 3 * h2d_bytes / t_copy / 1e9,
                                   "what": "same pinned source windows over 4 streams per GPU, no kernels, all ranks at once"}
        e2e["frac_of_copy_ceiling"] = e2e_value / ceiling

        # the drop-in case: the caller's buffers are ordinary heap memory (a Rust Vec, Python bytes, a RustBuffer);
        # the library stages them through its pinned ring
        if not args.no_pageable:
            page_in = np.empty((nb, in_bytes), np.uint8)
            page_in[:] = host
            page_out = np.empty((nb, out_bytes), np.uint8)

            # the staging memcpy is the bound here (one host thread moves ~9 GB/s): eight workers per GPU instead of four
            cfg_p = ll.CBatchConfig(1, 8, dev)

            def pageable_step():
                for i in range(nb):
                    cin[i] = ll.CImage(SRC_W, SRC_H, page_in[i].ctypes.data, in_bytes)
                    cout[i] = ll.CImage(0, 0, page_out[i].ctypes.data, out_bytes)
                failed = lib.slimg_b200_pipeline_batch_host(cin, nb, C.byref(opts), order, C.byref(cfg_p), cout, st)
                if failed:
                    raise SystemExit("pageable e2e batch failed: " + lib.slimg_b200_last_error().decode())

            p_steps = max(3, min(args.steps, 4))
            dt_p, step_ms_p = timed_e2e(pageable_step, p_steps)
            same = bool((page_out == np.ctypeslib.as_array(C.cast(pin_out, C.POINTER(C.c_uint8)), shape=(nb, out_bytes))).all())
            e2e["pageable"] = {"value": _shard.whole_job_throughput(nb * p_steps * MPX_PER_IMAGE, world, dt_p), "unit": "Mpx/s",
                               "steps": p_steps, "step_ms": step_ms_p, "host_buffers": "pageable (numpy heap arrays), staged through the library's pinned ring", "streams_per_device": 8,
                               "equals_pinned_result": same}
            e2e["pageable"]["frac_of_pinned"] = e2e["pageable"]["value"] / e2e_value
            del page_in, page_out

        # a folder of photos is not one geometry: eight source sizes in one batch through the same call (one plan per
        # geometry from the library's plan cache, images of different sizes interleaved on the worker streams)
        if not args.no_pageable:
            sizes = [(4000, 3000), (6000, 4000), (7680, 4320), (1024, 1024), (3000, 4000), (1920, 1080), (2048, 2048), (4000, 6000)]
            fit = ll.make_options(resize=ll.ResizeMode.FIT(800, 800))
            offs, o = [], 0
            for i in range(nb):
                w, h = sizes[i % len(sizes)]
                offs.append((o, w, h))
                o += w * h * 4
            assert o <= nb * in_bytes and 800 * 800 * 4 <= out_bytes
            mixed_px = sum(w * h for _, w, h in offs)

            def mixed_step():
                for i, (off, w, h) in enumerate(offs):
                    cin[i] = ll.CImage(w, h, pin_in + off, w * h * 4)
                    cout[i] = ll.CImage(0, 0, pin_out + i * out_bytes, out_bytes)
                if lib.slimg_b200_pipeline_batch_host(cin, nb, C.byref(fit), ll.ORDER_CONVERT, C.byref(cfg), cout, st):
                    raise SystemExit("mixed-size e2e batch failed: " + lib.slimg_b200_last_error().decode())

            m_steps = max(3, min(args.steps, 4))
            dt_m, step_ms_m = timed_e2e(mixed_step, m_steps)
            e2e["mixed_sizes"] = {"value": _shard.whole_job_throughput(m_steps * mixed_px / 1e6, world, dt_m), "unit": "Mpx/s",
                                  "what": "12 MP / 24 MP / 8K / 1024^2 / portrait / 1080p / 2048^2 photos in one batch, Fit(800, 800), pinned buffers",
                                  "images_per_gpu_per_step": nb, "geometries": len(sizes), "steps": m_steps, "step_ms": step_ms_m,
                                  "host_image_bytes_per_step": o}

        # the other extreme: no two images of the run share a geometry (user crops, scans): every image costs a plan on the
        # host (taps, fixed-point weights, strip layout: 5-8 ms for 12 MP) before its kernel can start; the scheduler's workers
        # build their plans concurrently, outside the cache lock
        if not args.no_pageable:
            uniq = {"k": 0}

            def unique_step():
                for i in range(nb):
                    w = 4000 - 4 * (uniq["k"] % 400)
                    uniq["k"] += 1
                    cin[i] = ll.CImage(w, 3000, pin_in + i * in_bytes, w * 3000 * 4)
                    cout[i] = ll.CImage(0, 0, pin_out + i * out_bytes, out_bytes)
                if lib.slimg_b200_pipeline_batch_host(cin, nb, C.byref(fit), ll.ORDER_CONVERT, C.byref(cfg), cout, st):
                    raise SystemExit("unique-size e2e batch failed: " + lib.slimg_b200_last_error().decode())

            k0 = None
            u_steps = 3
            for _ in range(2):
                unique_step()
            barrier()
            k0 = uniq["k"]
            t0 = time.perf_counter()
            px_u = 0
            for _ in range(u_steps):
                kk = uniq["k"]
                unique_step()
                px_u += sum((4000 - 4 * ((kk + i) % 400)) * 3000 for i in range(nb))
            torch.cuda.synchronize()
            dt_u = _shard.max_over_ranks(time.perf_counter() - t0, device="cuda")
            e2e["unique_sizes"] = {"value": _shard.whole_job_throughput(px_u / 1e6, world, dt_u), "unit": "Mpx/s",
                                   "what": "every image a geometry of its own (3000 rows, widths 4000, 3996, ... never repeated inside the plan cache's reach), Fit(800, 800), pinned buffers: one host-side plan per image",
                                   "images_per_gpu_per_step": nb, "steps": u_steps, "streams_per_device": 4}

        # the library's own multi-device scheduler: ONE process (rank 0) feeding every GPU of the box (n_devices = world),
        # no Python per rank -- the `--jobs` analogue proper; the other ranks wait
        if world > 1:
            barrier()
            if rank == 0:
                n_all = nb * world
                cin_all = (ll.CImage * n_all)()
                cout_all = (ll.CImage * n_all)()
                st_all = (C.c_int * n_all)()
                outs_all = lib.slimg_b200_host_alloc(n_all * out_bytes)
                cfg_all = ll.CBatchConfig(world, 4, None)

                def all_step():
                    for i in range(n_all):
                        cin_all[i] = ll.CImage(SRC_W, SRC_H, pin_in + (i % nb) * in_bytes, in_bytes)
                        cout_all[i] = ll.CImage(0, 0, outs_all + i * out_bytes, out_bytes)
                    if lib.slimg_b200_pipeline_batch_host(cin_all, n_all, C.byref(opts), order, C.byref(cfg_all), cout_all, st_all):
                        raise SystemExit("one-process e2e batch failed: " + lib.slimg_b200_last_error().decode())

                for _ in range(2):
                    all_step()
                t0 = time.perf_counter()
                for _ in range(3):
                    all_step()
                dt_all = time.perf_counter() - t0
                e2e["one_process_all_devices"] = {"value": n_all * 3 * MPX_PER_IMAGE / dt_all, "unit": "Mpx/s", "n_devices": world,
                                                  "streams_per_device": 4, "images_per_step": n_all}
                lib.slimg_b200_host_free(outs_all)
            barrier()
        lib.slimg_b200_host_free(pin_in); lib.slimg_b200_host_free(pin_out)

    configs = None
    if world == 1 and not args.no_configs:
        configs = config_table(ll, torch, peak, src)

    cpu = None
    if rank == 0 and not args.no_cpu:
        threads = host_threads()
        n_img = args.cpu_images or max(32, min(256, threads * 8))
        dt, mpx = cpu_run(args.order, n_img, threads)
        dt1, mpx1 = cpu_run(args.order, 3, 1)          # --jobs 1: one image at a time
        cpu = {"value": mpx, "unit": "Mpx/s", "cores": threads, "kind": "port", "jobs": threads,
               "sample": f"{n_img} synthetic 24 MP images of the same workload, one image per thread, {dt:.1f} s wall",
               "jobs_1": {"value": mpx1, "unit": "Mpx/s", "cores": 1, "jobs": 1,
                          "sample": f"3 images of the same workload on one thread, {dt1:.1f} s wall"}}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        _emit(json.dumps({
            "metric": "Mpx/s resize->crop->extend", "value": value, "unit": "Mpx/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": workload_config(args, B), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks, "configs": configs, "build_id": build_id(),
        }))


if __name__ == "__main__":
    main()
