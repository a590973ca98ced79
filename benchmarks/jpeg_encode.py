"""Throughput of this library's JPEG encoder (csrc/jpeg_encode.cu), Pillow's default settings (quality 75, 4:2:0):
32-frame batches of 640x480 (the resized frames of BASELINE config 1 / 2) and of 1080p; CUDA events around the
encode calls, then the same with several batches in flight; Pillow on one host core for scale."""
import io
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from aiko_services_b200 import jpeg  # noqa: E402


def main():
    from PIL import Image
    dev = torch.device("cuda:0")
    for name, (h, w) in (("640x480", (480, 640)), ("1080p", (1080, 1920))):
        host = np.stack([bench.natural_frame(900 + k, h, w) for k in range(32)])
        frames = torch.from_numpy(host).to(dev)
        enc = jpeg.GpuJpegEncoder(dev)
        files = enc.encode(frames)
        mean_kb = np.mean([len(f) for f in files]) / 1e3
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            enc.encode_async(frames)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"encoder, 32 x {name}: {ms:.3f} ms per batch on the device ({32 / ms * 1e3:.0f} frames/s), {mean_kb:.0f} KB per file")
        for lanes in (4, 8):
            streams = [torch.cuda.Stream(dev) for _ in range(lanes)]
            encs = [jpeg.GpuJpegEncoder(dev) for _ in range(lanes)]
            for k in range(lanes):
                encs[k].encode_async(frames, stream=streams[k])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rounds = 6
            for _ in range(rounds):
                for k in range(lanes):
                    encs[k].encode_async(frames, stream=streams[k])
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            print(f"encoder, 32 x {name}, {lanes} batches in flight: {32 * lanes * rounds / dt:.0f} frames/s")
        Image.fromarray(host[0], "RGB").save(io.BytesIO(), "JPEG")
        t0 = time.perf_counter()
        for k in range(32):
            Image.fromarray(host[k], "RGB").save(io.BytesIO(), "JPEG")
        print(f"Pillow, one core, {name}: {32 / (time.perf_counter() - t0):.0f} frames/s")


if __name__ == "__main__":
    main()
