"""One-off sweep of ImageOverlay's drawing over random frame sizes and random detection lists:
overlay alone (copy + cell kernel) and fused behind a resize, with one list shared by the frames or
one list per frame, against the CPU replay of the same draw list (tests/util.py) AND, where OpenCV is
installed, against the reference element's own cv2 calls (oracle/reference_port.py).
python benchmarks/stress_overlay.py [seed] [cases]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from aiko_services_b200 import ops  # noqa: E402
from aiko_services_b200.overlay_plan import build_ops  # noqa: E402
from aiko_services_b200.text_atlas import TextAtlas  # noqa: E402
from oracle import image as O  # noqa: E402
from oracle import reference_port  # noqa: E402
from aiko_services_b200.overlay_plan import plan  # noqa: E402
from util import overlay_cpu, random_overlay  # noqa: E402


def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    cases = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda:0")
    atlas = TextAtlas()
    bad = 0
    t0 = time.time()
    for i in range(cases):
        h, w = int(rng.integers(20, 300)), int(rng.integers(40, 500))
        n = int(rng.integers(1, 4))
        frames = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
        overlay = random_overlay(rng, w, h, int(rng.integers(1, 7)))
        d = torch.from_numpy(frames).to(dev)
        if i % 2 == 0:      # overlay alone on the frames as they are
            ov = ops.DeviceOverlay.from_host([build_ops(overlay, w, h, "rgb", atlas)] * n, dev, atlas.device(dev), width=w, height=h)
            got = ops.overlay_draw(d, ov).cpu().numpy()
            want = [overlay_cpu(f, overlay, atlas) for f in frames]
        else:               # resize -> overlay in one call
            W, H = max(4, int(rng.integers(8, 300)) & ~3), int(rng.integers(8, 200))
            overlay = random_overlay(rng, W, H, int(rng.integers(1, 7)))
            ov = ops.DeviceOverlay.from_host([build_ops(overlay, W, H, "rgb", atlas)] * n, dev, atlas.device(dev), width=W, height=H)
            got = ops.resize_linear(d, W, H, overlay=ov).cpu().numpy()
            want = [overlay_cpu(O.resize_linear(f, W, H), overlay, atlas) for f in frames]
        if i % 3 == 0:      # the same through the planner's packed buffer, one list per frame, vs cv2 itself
            try:
                import cv2  # noqa: F401
                shape = (w, h) if i % 2 == 0 else (W, H)
                per_frame = [random_overlay(rng, shape[0], shape[1], int(rng.integers(1, 6))) for _ in range(n)]
                ovp = ops.DeviceOverlay.from_plan(plan(per_frame, shape[0], shape[1], "rgb"), dev)
                if i % 2 == 0:
                    got = ops.overlay_draw(d, ovp).cpu().numpy()
                    want = [reference_port.image_overlay([f], o)[0] for f, o in zip(frames, per_frame)]
                else:
                    got = ops.resize_linear(d, W, H, overlay=ovp).cpu().numpy()
                    want = [reference_port.image_overlay([cv2.resize(f, (W, H))], o)[0] for f, o in zip(frames, per_frame)]
            except ModuleNotFoundError:
                pass
        for k in range(n):
            if not np.array_equal(got[k], want[k]):
                bad += 1
                print("MISMATCH case", i, frames.shape, "differing bytes", int((got[k] != want[k]).sum()))
                break
    print(f"{cases} cases, {bad} mismatches, {time.time() - t0:.1f} s")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
