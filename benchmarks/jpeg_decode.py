"""Throughput of this library's JPEG decoder (csrc/jpeg_decode.cu) on 32 x 1080p frames, by restart interval,
next to nvJPEG's hybrid decoder; kernel times from CUDA events around the whole decode call."""
import io
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from aiko_services_b200 import jpeg  # noqa: E402


def blob(seed, **kw):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(bench.natural_frame(seed), "RGB").save(buf, "JPEG", quality=90, subsampling=2, **kw)
    return buf.getvalue()


def pinned(b):
    t = torch.empty(len(b), dtype=torch.uint8).pin_memory()
    t.numpy()[...] = np.frombuffer(b, np.uint8)
    return t


def main():
    dev = torch.device("cuda:0")
    frames = torch.empty((32, 1080, 1920, 3), dtype=torch.uint8, device=dev)
    for label, kw in (("no restart markers", {}), ("restart every MCU row (120 MCUs)", {"restart_marker_rows": 1}),
                      ("restart every 30 MCUs", {"restart_marker_blocks": 30}),
                      ("restart every 8 MCUs", {"restart_marker_blocks": 8})):
        blobs = [pinned(blob(500 + k, **kw)) for k in range(32)]
        dec = jpeg.GpuJpegDecoder(dev)
        dec.decode(blobs, out=frames, check=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(10):
            dec.decode(blobs, out=frames)
        e1.record()
        host = (time.perf_counter() - t0) / 10 * 1e3
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"own decoder, {label}: {ms:.2f} ms per 32 frames on the device ({32 / ms * 1e3:.0f} frames/s), "
              f"host {host:.2f} ms per call, {np.mean([b.numel() for b in blobs]) / 1e3:.0f} KB per file")
    # several batches in flight on separate CUDA streams: the entropy kernel of one 32-image batch keeps only
    # 68 warps busy (one thread per restart segment), so concurrent batches fill the rest of the GPU
    blobs = [pinned(blob(500 + k)) for k in range(32)]      # Pillow's default output: no restart markers
    for stages, what in ((15, "all stages"), (3, "segments + entropy decode only"), (12, "IDCT + colour only")):
        os.environ["AIKO_B200_JPEG_STAGES"] = str(stages)
        for lanes in (1, 4, 8, 12, 16) if stages != 12 else (1, 8):
            streams = [torch.cuda.Stream(dev) for _ in range(lanes)]
            decs = [jpeg.GpuJpegDecoder(dev) for _ in range(lanes)]
            outs = [torch.empty((32, 1080, 1920, 3), dtype=torch.uint8, device=dev) for _ in range(lanes)]
            for k in range(lanes):
                decs[k].decode(blobs, out=outs[k], stream=streams[k])
            torch.cuda.synchronize()
            rounds = 4
            call_s = 0.0
            t0 = time.perf_counter()
            for r in range(rounds):
                for k in range(lanes):
                    if decs[k]._event is not None:
                        decs[k]._event.synchronize()
                    c0 = time.perf_counter()
                    decs[k].decode(blobs, out=outs[k], stream=streams[k])
                    call_s += time.perf_counter() - c0
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            print(f"own decoder, no restart markers, {what}, {lanes} batches in flight: "
                  f"{32 * lanes * rounds / dt:.0f} frames/s, host {call_s / (lanes * rounds) * 1e3:.2f} ms per call")
            del outs, decs
    os.environ.pop("AIKO_B200_JPEG_STAGES")
    blobs = [pinned(blob(500 + k)) for k in range(32)]
    nv = jpeg.JpegDecoder(dev, 32)
    nv.decode(blobs, out=frames)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        nv.decode(blobs, out=frames)
    torch.cuda.synchronize()
    print(f"nvJPEG hybrid: {160 / (time.perf_counter() - t0):.0f} frames/s")


if __name__ == "__main__":
    main()
