"""Launch one op a few times on its BASELINE-sized input: the target of `ncu -k <kernel>` captures
and of quick A/B timings (`--time`), e.g.

    python benchmarks/one_kernel.py bicubic --time
    ncu --set full --clock-control none --import-source on -k regex:bicubic_mma -c 2 -o gpurun_out/bicubic \
        python benchmarks/one_kernel.py bicubic
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aiko_services_b200 import audio_ops, ops  # noqa: E402


def make(name, dev):
    g = torch.Generator(device=dev).manual_seed(1)
    rot = 4
    if name in ("bicubic", "linear", "flip", "swap_rb", "bgr2gray"):
        hd = [torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        if name == "bicubic":
            out = torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev)
            return lambda i: ops.resize_bicubic_aa(hd[i % rot], 640, 480, out=out), 32
        if name == "linear":
            out = torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev)
            return lambda i: ops.resize_linear(hd[i % rot], 640, 480, out=out), 32
        if name == "bgr2gray":
            gray = torch.empty((32, 1080, 1920), dtype=torch.uint8, device=dev)
            return lambda i: ops.bgr2gray(hd[i % rot], out=gray), 32
        out = torch.empty_like(hd[0])
        if name == "flip":
            return lambda i: ops.flip(hd[i % rot], 1, out=out), 32
        return lambda i: ops.swap_rb(hd[i % rot], out=out), 32
    if name in ("step", "step_draw"):      # the bench.py step: 1080p -> 640x480 + swap, then the draw list in place
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        import bench
        from aiko_services_b200.overlay_plan import plan
        hd = [torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        out = torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev)
        ov = ops.DeviceOverlay.from_plan(plan(bench.OVERLAY, 640, 480, "bgr", n_frames=32), dev)
        if name == "step":
            return lambda i: ops.resize_linear(hd[i % rot], 640, 480, swap_rb=True, overlay=ov, out=out), 32
        ops.resize_linear(hd[0], 640, 480, swap_rb=True, out=out)
        return lambda i: ops.overlay_draw(out, ov, out=out), 32
    if name == "overlay":                  # ImageOverlay alone on 32 x 1080p frames (copy + draw)
        from aiko_services_b200.overlay_plan import build_ops
        from aiko_services_b200.text_atlas import TextAtlas
        atlas = TextAtlas.shared()
        overlay = {"rectangles": [{"x": 100 + 400 * i, "y": 150 + 180 * i, "w": 300, "h": 250} for i in range(4)],
                   "objects": [{"name": n, "confidence": 0.9 - 0.1 * i} for i, n in enumerate(("person", "car", "dog", "bicycle"))]}
        per_frame = [build_ops(overlay, 1920, 1080, "rgb", atlas)] * 32
        ov = ops.DeviceOverlay.from_host(per_frame, dev, atlas.device(dev), width=1920, height=1080)
        hd = [torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        out = torch.empty_like(hd[0])
        return lambda i: ops.overlay_draw(hd[i % rot], ov, out=out), 32
    if name == "uhd":                      # 4K -> 1080p, 8 frames (199 MB in, 50 MB out)
        src = [torch.randint(0, 256, (8, 2160, 3840, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        out = torch.empty((8, 1080, 1920, 3), dtype=torch.uint8, device=dev)
        return lambda i: ops.resize_linear(src[i % rot], 1920, 1080, out=out), 8
    if name in ("vga", "gray"):
        shape = (256, 480, 640, 3) if name == "vga" else (32, 1080, 1920)
        size = (320, 240) if name == "vga" else (640, 480)
        src = [torch.randint(0, 256, shape, dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        out = torch.empty((shape[0], size[1], size[0]) + shape[3:], dtype=torch.uint8, device=dev)
        return lambda i: ops.resize_linear(src[i % rot], size[0], size[1], out=out), shape[0]
    if name == "boxes":                    # ImageOverlay planned and drawn on the GPU from device-resident detections
        import bench
        frames = [torch.randint(0, 256, (32, 480, 640, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        rects, objs = bench.OVERLAY["rectangles"], bench.OVERLAY["objects"]
        names = sorted({o["name"] for o in objs})
        det = ops.DeviceDetections(
            torch.tensor([[[r["x"], r["y"], r["w"], r["h"]] for r in rects]], dtype=torch.int32, device=dev),
            torch.tensor([[names.index(o["name"]) for o in objs]], dtype=torch.int32, device=dev),
            torch.tensor([[o["confidence"] for o in objs]], dtype=torch.float32, device=dev), names)
        assert int(ops.overlay_boxes(frames[0], det, layout="bgr", check=False)[1].cpu()[0]) == 0
        return lambda i: ops.overlay_boxes(frames[i % rot], det, layout="bgr", check=False), 32
    if name == "alpha":                    # a frame-sized RGBA image blended over 32 x 640x480 frames, in place
        frames = [torch.randint(0, 256, (32, 480, 640, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        rgba = torch.randint(0, 256, (480, 640, 4), dtype=torch.uint8, device=dev, generator=g)
        return lambda i: ops.alpha_paste(frames[i % rot], rgba), 32
    if name == "norm":                     # BASELINE config 4: 4K -> 224x224 fp32 NCHW, normalised
        src = [torch.randint(0, 256, (32, 2160, 3840, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(2)]
        out = torch.empty((32, 3, 224, 224), dtype=torch.float32, device=dev)
        return lambda i: ops.resize_norm_f32(src[i % 2], 224, 224, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225), out=out), 32
    if name == "jpeg":                     # 32 x 1080p JPEG files as Pillow writes them (no restart markers)
        import bench
        from aiko_services_b200 import jpeg
        blobs = []
        for k in range(32):
            b = bench.jpeg_blob(500 + k)
            t = torch.empty(len(b), dtype=torch.uint8).pin_memory()
            t.numpy()[...] = np.frombuffer(b, np.uint8)
            blobs.append(t)
        dec = jpeg.GpuJpegDecoder(dev)
        frames = torch.empty((32, 1080, 1920, 3), dtype=torch.uint8, device=dev)
        return lambda i: dec.decode(blobs, out=frames), 32
    if name == "jpegenc":                  # 32 x 640x480 frames -> JPEG files (Pillow's default settings)
        import bench
        from aiko_services_b200 import jpeg
        host = np.stack([bench.natural_frame(900 + k, 480, 640) for k in range(32)])
        frames = torch.from_numpy(host).to(dev)
        enc = jpeg.GpuJpegEncoder(dev)
        return lambda i: enc.encode_async(frames), 32
    if name in ("audio", "fft"):
        t = np.arange(16000) / 16000.0
        freqs = np.geomspace(50, 7000, 1024)
        rng = np.random.default_rng(5000)
        win = (3000 * np.sin(2 * np.pi * freqs[:, None] * t[None, :]) + rng.normal(0, 200, (1024, 16000))).astype(np.int16)
        aud = [torch.from_numpy(np.roll(win, k, axis=0)).to(dev) for k in range(rot)]
        if name == "audio":
            return lambda i: audio_ops.audio_chain(aud[i % rot], 16000), 1024
        return lambda i: audio_ops.audio_chain(aud[i % rot], 16000, bands=False, top=False), 1024
    raise SystemExit(f"unknown op {name}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("op")
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--steps", type=int, default=50)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    fn, units = make(args.op, dev)
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    if args.time:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / args.steps * 1e3
        tag = " ".join(f"{k}={v}" for k, v in sorted(os.environ.items()) if k.startswith("AIKO_B200_"))
        print(f"{args.op}: {us:.1f} us per launch, {units / us * 1e6:.0f} units/s {tag}")


if __name__ == "__main__":
    main()
