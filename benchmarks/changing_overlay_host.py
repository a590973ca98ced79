"""Host-side cost of the changing-overlay step of bench.py, split into its parts (per step, us):
planning (ring.plan), the drawing call (resize + overlay enqueue), and where the slow steps are."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from aiko_services_b200 import ops  # noqa: E402
from aiko_services_b200.text_atlas import TextAtlas  # noqa: E402


def changing_overlay(i):
    o = bench.OVERLAY
    return {"rectangles": [dict(r, x=r["x"] + (i * 7) % 23, y=r["y"] + (i * 5) % 17) for r in o["rectangles"]],
            "objects": [dict(ob, confidence=0.01 + ((i * 13 + 29 * k) % 99) / 100.0) for k, ob in enumerate(o["objects"])]}


def main():
    dev = torch.device("cuda:0")
    ring = ops.OverlayRing(dev)
    src = [torch.zeros((32, 1080, 1920, 3), dtype=torch.uint8, device=dev) for _ in range(4)]
    out = torch.empty((32, 480, 640, 3), dtype=torch.uint8, device=dev)
    dicts = [changing_overlay(i) for i in range(403)]
    for lap in range(2):
        torch.cuda.synchronize()
        tp, tr = [], []
        v0 = TextAtlas.shared().view().n_patches
        t_all = time.perf_counter()
        for i in range(200):
            t0 = time.perf_counter()
            ov = ring.plan(dicts[lap * 200 + i], 640, 480, "bgr", n_frames=32)
            t1 = time.perf_counter()
            ops.resize_linear(src[i % 4], 640, 480, swap_rb=True, overlay=ov, out=out)
            t2 = time.perf_counter()
            tp.append((t1 - t0) * 1e6)
            tr.append((t2 - t1) * 1e6)
        host = (time.perf_counter() - t_all) / 200 * 1e6
        torch.cuda.synchronize()
        total = (time.perf_counter() - t_all) / 200 * 1e6
        tp, tr = np.array(tp), np.array(tr)
        print(f"lap {lap}: host {host:.1f} us/step, with drain {total:.1f}; plan median {np.median(tp):.1f} mean {tp.mean():.1f} "
              f"max {tp.max():.0f}; draw call median {np.median(tr):.1f} mean {tr.mean():.1f} max {tr.max():.0f}; "
              f"new patches {TextAtlas.shared().view().n_patches - v0}; slow plan steps {np.nonzero(tp > 100)[0][:12].tolist()}")


if __name__ == "__main__":
    main()
