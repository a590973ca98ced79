"""Random sweep of the device-side overlay planner (ops.overlay_boxes) against the host planner: frame sizes, box
positions (also outside the frame), confidences (incl. rounding ties and values >= 1), unnamed classes, shared and
per-frame lists.  Prints the number of mismatches (expected: 0) and how often the host path was needed."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aiko_services_b200 import ops  # noqa: E402
from aiko_services_b200.overlay_plan import plan  # noqa: E402
from aiko_services_b200.text_atlas import TextAtlas  # noqa: E402


def main():
    trials = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    dev = torch.device("cuda:0")
    atlas = TextAtlas.shared()
    rng = np.random.default_rng(404)
    names = ["person", "bicycle", "car", "", "traffic light", "Zq{|}~", "x", "a much longer class name, 39 chars.."]
    bad = fallback = 0
    for t in range(trials):
        W, H = int(rng.integers(64, 900)), int(rng.integers(48, 700))
        n = int(rng.integers(1, 4))
        lists = 1 if rng.random() < 0.5 else n
        slots = int(rng.integers(1, 12))
        boxes = np.stack([rng.integers(-40, W + 40, (lists, slots)), rng.integers(-40, H + 40, (lists, slots)),
                          rng.integers(-10, W, (lists, slots)), rng.integers(-10, H, (lists, slots))], -1).astype(np.int32)
        if t % 3 and W > 420 and H > 200:          # two trials of three: boxes whose labels stay clear of the borders
            boxes[..., 0] = rng.integers(10, W - 400, (lists, slots))
            boxes[..., 1] = rng.integers(40, H - 110, (lists, slots))
            boxes[..., 2] = rng.integers(0, 60, (lists, slots))
            boxes[..., 3] = rng.integers(0, 60, (lists, slots))
        cls = rng.integers(-2, len(names) + 1, (lists, slots)).astype(np.int32)
        conf = rng.random((lists, slots)).astype(np.float32)
        mode = rng.random((lists, slots))
        conf[mode < 0.1] = (rng.integers(0, 200, int((mode < 0.1).sum())) / 200.0 + 0.0025).astype(np.float32)   # near ties
        conf[(mode >= 0.1) & (mode < 0.15)] = rng.choice([0.0, -1.0, 1.0, 7.5, 123.456, 0.005, 0.995], int(((mode >= 0.1) & (mode < 0.15)).sum()))
        frames = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
        det = ops.DeviceDetections(torch.from_numpy(boxes).to(dev), torch.from_numpy(cls).to(dev), torch.from_numpy(conf).to(dev), names)
        dicts = det.to_overlay_dicts()
        layout = "rgb" if t % 2 else "bgr"
        d = torch.from_numpy(frames).to(dev)
        got, status = ops.overlay_boxes(d, det, layout=layout, out=torch.empty_like(d), check=False)
        code = int(status.cpu()[0])
        planned = plan(dicts[0] if lists == 1 else dicts, W, H, layout, n_frames=n)
        want = ops.overlay_draw(d, ops.DeviceOverlay.from_plan(planned, dev, atlas.device(dev))).cpu().numpy()
        if code == 0:
            if not np.array_equal(got.cpu().numpy(), want):
                bad += 1
                print(f"mismatch: trial {t}, {W}x{H}, {lists} lists x {slots} slots, layout {layout}")
        else:
            fallback += 1
            if not np.array_equal(got.cpu().numpy(), frames):
                bad += 1
                print(f"drew although status {code}: trial {t}")
            checked = ops.overlay_boxes(d, det, layout=layout, out=torch.empty_like(d))
            if not np.array_equal(checked.cpu().numpy(), want):
                bad += 1
                print(f"fallback mismatch: trial {t}")
    print(f"stress_boxes: {trials} trials, {bad} mismatches, host path needed in {fallback}")


if __name__ == "__main__":
    main()
