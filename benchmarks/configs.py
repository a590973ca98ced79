#!/usr/bin/env python
"""Secondary measurements: the other BASELINE.json configs and the stand-alone kernels, each as
units/s and as a fraction of the measured HBM roofline (algorithmic bytes of SURVEY.md 8(d)).
Device-resident inputs, CUDA events, rotating inputs larger than L2.  One JSON line per kernel.

    python benchmarks/configs.py [--steps 50]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aiko_services_b200 import audio_ops, ops  # noqa: E402


def peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return (float(json.load(open(path))["hbm_gbs"]), "measured") if os.path.exists(path) else (6650.0, "fallback")


def timed(fn, steps, warmup=5):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def report(name, units, unit_name, algo_bytes_per_unit, ms, extra=None):
    peak, src = peak_gbs()
    achieved = units * algo_bytes_per_unit / (ms * 1e-3) / 1e9
    line = {"kernel": name, "value": units / (ms * 1e-3), "unit": f"{unit_name}/s", "ms_per_launch": ms,
            "units_per_launch": units, "algorithmic_bytes_per_unit": algo_bytes_per_unit,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": src}}
    line.update(extra or {})
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    rot = 4

    hd = [torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
    out = torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev)
    ms = timed(lambda i: ops.resize_linear(hd[i % rot], 640, 480, out=out), args.steps)
    report("resize_linear 1080p->640x480 (no overlay)", 32, "frames", 6_451_200, ms)
    ms = timed(lambda i: ops.resize_bicubic_aa(hd[i % rot], 640, 480, out=out), max(args.steps // 5, 5))
    report("resize_bicubic_aa 1080p->640x480 (config 1, codecs excluded)", 32, "frames", 7_142_400, ms)

    # the generic row kernel (gray, RGBA), and RGB rows that are not 16-byte aligned (row pipeline, per-row offsets)
    gray_in = [torch.randint(0, 256, (32, 1080, 1920), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
    gray_out = torch.empty((32, 480, 640), dtype=torch.uint8, device=dev)
    ms = timed(lambda i: ops.resize_linear(gray_in[i % rot], 640, 480, out=gray_out), args.steps)
    report("resize_linear gray 1080p->640x480", 32, "frames", 960 * 1920 + 640 * 480, ms)
    del gray_in, gray_out
    odd_in = [torch.randint(0, 256, (32, 1080, 1918, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
    ms = timed(lambda i: ops.resize_linear(odd_in[i % rot], 640, 480, out=out), args.steps)
    report("resize_linear 1918x1080->640x480 (rows off 16-byte boundaries)", 32, "frames", 960 * 1918 * 3 + 921_600, ms)
    del odd_in

    vga = [torch.randint(0, 256, (256, 480, 640, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
    vga_out = torch.empty((256, 240, 320, 3), dtype=torch.uint8, device=dev)
    ms = timed(lambda i: ops.resize_linear(vga[i % rot], 320, 240, out=vga_out), args.steps)
    report("resize_linear 640x480->320x240 (short rows: several per ring slot)", 256, "frames", 640 * 480 * 3 + 320 * 240 * 3, ms)
    del vga, vga_out

    sw = torch.empty_like(hd[0])
    ms = timed(lambda i: sw.copy_(hd[i % rot]), args.steps)
    report("calibration: torch device copy of one 199 MB batch (read + write)", 32, "frames", 2 * 6_220_800, ms)
    ms = timed(lambda i: ops.swap_rb(hd[i % rot], out=sw), args.steps)
    report("swap_rb 1080p", 32, "frames", 2 * 6_220_800, ms)
    gray = torch.empty((32, 1080, 1920), dtype=torch.uint8, device=dev)
    ms = timed(lambda i: ops.bgr2gray(hd[i % rot], out=gray), args.steps)
    report("bgr2gray 1080p", 32, "frames", 6_220_800 + 2_073_600, ms)
    ms = timed(lambda i: ops.flip(hd[i % rot], 1, out=sw), args.steps)
    report("flip horizontal 1080p", 32, "frames", 2 * 6_220_800, ms)
    del sw, gray

    uhd = [torch.randint(0, 256, (32, 2160, 3840, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
    nchw = torch.empty((32, 3, 224, 224), dtype=torch.float32, device=dev)
    ms = timed(lambda i: ops.resize_norm_f32(uhd[i % rot], 224, 224, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225),
                                             out=nchw), args.steps)
    report("resize_norm_f32 4K->224x224 NCHW (config 4, batch 32)", 32, "frames", 4_272_128, ms)
    del uhd

    # round 2: YUV, alpha blending, and the JPEG codec (algorithmic bytes: compressed file + RGB frame)
    yuv = torch.empty_like(hd[0])
    ms = timed(lambda i: ops.yuv(hd[i % rot], out=yuv), args.steps)
    report("yuv (RGB2YUV) 1080p", 32, "frames", 2 * 6_220_800, ms)
    del yuv
    small = [torch.randint(0, 256, (32, 480, 640, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
    rgba = torch.randint(0, 256, (480, 640, 4), dtype=torch.uint8, device=dev, generator=g)
    ms = timed(lambda i: ops.alpha_paste(small[i % rot], rgba), args.steps)
    report("alpha_paste, frame-sized RGBA image over 640x480 frames, in place", 32, "frames", 2 * 921_600, ms)
    sys.path.insert(0, ROOT)
    import bench
    from aiko_services_b200 import jpeg
    blobs = []
    for k in range(32):
        b = bench.jpeg_blob(500 + k)
        pinned = torch.empty(len(b), dtype=torch.uint8).pin_memory()
        pinned.numpy()[...] = np.frombuffer(b, np.uint8)
        blobs.append(pinned)
    file_bytes = int(np.mean([b.numel() for b in blobs]))
    dec, enc = jpeg.GpuJpegDecoder(dev), jpeg.GpuJpegEncoder(dev)
    frames = torch.empty((32, 1080, 1920, 3), dtype=torch.uint8, device=dev)
    ms = timed(lambda i: dec.decode(blobs, out=frames), max(args.steps // 5, 5))
    report("jpeg decode 1080p 4:2:0 quality 90, no restart markers (one batch at a time, H2D of the files included)", 32,
           "frames", file_bytes + 6_220_800, ms)
    natural = torch.from_numpy(np.stack([bench.natural_frame(900 + k, 480, 640) for k in range(32)])).to(dev)
    ms = timed(lambda i: enc.encode_async(natural), args.steps)
    report("jpeg encode 640x480, quality 75 4:2:0 (one batch at a time)", 32, "frames", 921_600 + 44_000, ms)
    del frames, small

    t = np.arange(16000) / 16000.0
    freqs = np.geomspace(50, 7000, 1024)
    rng = np.random.default_rng(5000)
    win = (3000 * np.sin(2 * np.pi * freqs[:, None] * t[None, :]) + rng.normal(0, 200, (1024, 16000))).astype(np.int16)
    aud = [torch.from_numpy(np.roll(win, k, axis=0)).to(dev) for k in range(rot)]
    ms = timed(lambda i: audio_ops.audio_chain(aud[i % rot], 16000), args.steps)
    report("audio_chain 1024 x 16000 int16 (config 5: FFT + bands + gate/top-100)", 1024, "windows", 96_832, ms)
    ms = timed(lambda i: audio_ops.audio_chain(aud[i % rot], 16000, bands=False, top=False), args.steps)
    report("audio_chain, FFT magnitudes only", 1024, "windows", 96_000, ms)


if __name__ == "__main__":
    main()
