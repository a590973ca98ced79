"""Does the overlay pass of batch i overlap the resize of batch i+1 when consecutive batches alternate
between two CUDA streams (two output buffers)?  Prints us per step for 1 and 2 streams."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from aiko_services_b200 import ops  # noqa: E402
from aiko_services_b200.overlay_plan import plan  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    src = [torch.randint(0, 256, (32, 1080, 1920, 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(4)]
    ov = ops.DeviceOverlay.from_plan(plan(bench.OVERLAY, 640, 480, "bgr", n_frames=32), dev)
    for lanes in (1, 2, 3):
        streams = [torch.cuda.Stream(dev) for _ in range(lanes)]
        outs = [torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev) for _ in range(lanes)]
        steps = 300

        def run(n):
            for i in range(n):
                s = streams[i % lanes]
                ops.resize_linear(src[i % 4], 640, 480, swap_rb=True, overlay=ov, out=outs[i % lanes], stream=s)
        run(12)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        main_stream = torch.cuda.current_stream()
        e0.record(main_stream)
        for s in streams:
            s.wait_event(e0)
        run(steps)
        for s in streams:
            done = torch.cuda.Event()
            done.record(s)
            main_stream.wait_event(done)
        e1.record(main_stream)
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / steps * 1e3
        print(f"{lanes} stream(s): {us:.1f} us per step, {32 / us * 1e6:.0f} frames/s")


if __name__ == "__main__":
    main()
