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
  cpu_baseline  the CPU oracle (restated reference path) on this box's host cores, bounded sample

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


def measured_traffic(order: str, batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of the K2 launch from the committed ncu --set full capture
    (profiles/traffic.json: bytes per image of that capture), scaled to this launch's batch; None if absent."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[order]
        return float(t["dram_bytes_per_image"]) * batch
    except Exception:
        return None


# ---- clocks ------------------------------------------------------------------------------------------

class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
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

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, reasons = [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); out["sm_max_mhz"] = float(f[2])
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm)); out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
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

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ll.kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = ll.kernel_launches() - launches0
    ms_max = _shard.max_over_ranks(ms, device="cuda")
    clocks = sampler.stop() if rank == 0 else None
    value = _shard.whole_job_throughput(B * args.steps * MPX_PER_IMAGE, world, ms_max * 1e-3)

    # roofline of the (single) kernel of the step, from this rank's own launches
    peak, peak_kind = measured_peak_gbs()
    alg_bytes = algorithmic_bytes_per_image(args.order) * B
    launch_ms = ms / max(1, launches)
    achieved = alg_bytes / (launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": measured_traffic(args.order, B), "peak_kind": peak_kind, "kernel": "resize pipeline (K2)",
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
        for _ in range(3):                    # warm-up: stream-ordered pool growth, plan upload, page pinning
            e2e_step()
        barrier()
        step_ms = []
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            t1 = time.perf_counter()
            e2e_step()
            step_ms.append(round(1e3 * (time.perf_counter() - t1), 2))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        e2e_value = _shard.whole_job_throughput(nb * e2e_steps * MPX_PER_IMAGE, world, _shard.max_over_ranks(dt, device="cuda"))
        checksum = int(np.ctypeslib.as_array(C.cast(pin_out, C.POINTER(C.c_uint8)), shape=(nb * out_bytes,))[::4097].sum())
        e2e = {"value": e2e_value, "unit": "Mpx/s", "h2d_bytes_per_step": nb * h2d_bytes, "d2h_bytes_per_step": nb * out_bytes,
               "host_image_bytes_per_step": nb * in_bytes,
               "images_per_gpu_per_step": nb, "steps": e2e_steps, "timer": "host wall clock around the C-ABI calls, max over ranks",
               "result_checksum": checksum, "step_ms": step_ms}
        lib.slimg_b200_host_free(pin_in); lib.slimg_b200_host_free(pin_out)

    cpu = None
    if rank == 0 and not args.no_cpu:
        threads = host_threads()
        n_img = args.cpu_images or max(32, min(256, threads * 8))
        dt, mpx = cpu_run(args.order, n_img, threads)
        cpu = {"value": mpx, "unit": "Mpx/s", "cores": threads, "kind": "port",
               "sample": f"{n_img} synthetic 24 MP images of the same workload, one image per thread, {dt:.1f} s wall"}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        _emit(json.dumps({
            "metric": "Mpx/s resize->crop->extend", "value": value, "unit": "Mpx/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": workload_config(args, B), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks,
        }))


if __name__ == "__main__":
    main()
