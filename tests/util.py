"""Helpers shared by the tests (CPU side)."""
import hashlib

import numpy as np

from oracle import image as O


def check_golden(golden, key, arr):
    """Compare ``arr`` with the stored golden output (whole array, or hash + crops)."""
    arr = np.ascontiguousarray(arr)
    assert tuple(golden[key + "/shape"]) == arr.shape, key
    if key + "/full" in golden:
        np.testing.assert_array_equal(arr, golden[key + "/full"], err_msg=key)
    else:
        np.testing.assert_array_equal(arr[:32, :48], golden[key + "/crop"], err_msg=key)
        np.testing.assert_array_equal(arr[-16:, -24:], golden[key + "/crop_end"], err_msg=key)
    assert hashlib.sha256(arr.tobytes()).hexdigest() == str(golden[key + "/sha256"]), key


def replay_ops_cpu(img, ops, atlas):
    """Apply a draw list (aiko_services_b200.overlay_plan.build_ops) with the C oracle.

    img: uint8 [H, W, 3] in the layout the ops were built for; modified in place.
    """
    for op in ops:
        if op.kind == 0:
            colour = (op.p0 & 0xff, (op.p0 >> 8) & 0xff, (op.p0 >> 16) & 0xff)
            O.rectangle(img, op.x0, op.y0, op.x1, op.y1, colour)
        else:
            px, py = op.x0 - (op.p2 & 0xffff), op.y0 - (op.p2 >> 16)
            pids = (op.p0 & 0xffff, op.p0 >> 16, op.p1 & 0xffff)
            O.text_lut(img, px, py, [atlas.patch_ids[p] for p in pids], [atlas.luts] * 3)
    return img


def overlay_cpu(image, overlay, atlas):
    """ImageOverlay on one frame via build_ops + oracle replay (RGB or gray ndarray)."""
    from aiko_services_b200.overlay_plan import build_ops
    h, w = image.shape[:2]
    if image.ndim == 2:
        ops = build_ops(overlay, w, h, "bgr", atlas)
        bgr = replay_ops_cpu(O.gray2bgr(image), ops, atlas)
        return O.bgr2gray(bgr)
    ops = build_ops(overlay, w, h, "rgb", atlas)
    return replay_ops_cpu(image.copy(), ops, atlas)


def random_overlay(rng, w, h, n=4):
    names = ["person", "bicycle", "car", "dog", "traffic light", "tv", "Zq{|}~", "x"]
    objs, rects = [], []
    for _ in range(n):
        rw, rh = int(rng.integers(0, w)), int(rng.integers(0, h))
        rects.append({"x": int(rng.integers(-20, w)), "y": int(rng.integers(-20, h)), "w": rw, "h": rh})
        o = {"confidence": float(rng.random())}
        if rng.random() < 0.8:
            o["name"] = names[int(rng.integers(len(names)))]
        objs.append(o)
    return {"objects": objs, "rectangles": rects}
