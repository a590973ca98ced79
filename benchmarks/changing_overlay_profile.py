"""cProfile of OverlayRing.resize_draw (the changing-overlay step of bench.py): where the host time of one step goes."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aiko_services_b200 import ops  # noqa: E402
from changing_overlay_host import changing_overlay  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    lanes = 3
    rings = [ops.OverlayRing(dev) for _ in range(lanes)]
    streams = [torch.cuda.Stream(dev) for _ in range(lanes)]
    if os.environ.get("NATURAL"):
        import bench
        src = [torch.from_numpy(bench.synthetic_batch(i)).to(dev) for i in range(4)]
    else:
        src = [torch.zeros((32, 1080, 1920, 3), dtype=torch.uint8, device=dev) for _ in range(4)]
    outs = [torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev) for _ in range(lanes)]
    dicts = [changing_overlay(i) for i in range(2000)]

    def run(n):
        for i in range(n):
            k = i % lanes
            rings[k].resize_draw(src[i % 4], dicts[i], 640, 480, swap_rb=True, out=outs[k], stream=streams[k])

    run(300)
    torch.cuda.synchronize()
    for lap, n in enumerate((2000, 2000, 200, 200)):
        t0 = time.perf_counter()
        run(n)
        host = (time.perf_counter() - t0) / n * 1e6
        torch.cuda.synchronize()
        print(f"lap {lap}: {n} steps, host {host:.1f} us/step, with drain {(time.perf_counter() - t0) / n * 1e6:.1f}")
    prof = cProfile.Profile()
    prof.enable()
    run(2000)
    prof.disable()
    torch.cuda.synchronize()
    pstats.Stats(prof).sort_stats("tottime").print_stats(14)


if __name__ == "__main__":
    main()
