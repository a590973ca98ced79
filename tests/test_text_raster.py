"""CPU: the library's own rasteriser of cv2.putText(FONT_HERSHEY_SIMPLEX, 0.75, 2, LINE_AA)
(csrc/text_raster.hpp) and the glyph atlas it builds (csrc/overlay_plan.cpp).

Oracles: tests/golden/hershey_atlas_cv2.npz -- per-pixel transfer functions measured by rendering
with cv2 on all 256 backgrounds (tests/golden/make_atlas_golden.py) -- and, where cv2 is installed,
cv2.putText itself on random backgrounds and positions, border-clipped labels included."""
import os

import numpy as np
import pytest

from aiko_services_b200.overlay_plan import build_ops
from aiko_services_b200.text_atlas import PAD_X, PAD_Y, TextAtlas
from util import replay_ops_cpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def atlas():
    return TextAtlas.shared()


def test_interior_atlas_equals_the_cv2_measured_transfer_functions(atlas):
    gold = np.load(os.path.join(HERE, "golden", "hershey_atlas_cv2.npz"))
    gids, gluts = gold["patch_ids"], gold["luts"]
    ids, luts = atlas.patch_ids, atlas.luts
    n = gids.shape[0]
    assert ids.shape[0] >= n and ids.shape[1:] == gids.shape[1:]
    # same transfer function at every pixel of every (character, phase, component) patch
    np.testing.assert_array_equal(luts[ids[:n]], gluts[gids])
    assert (luts[0] == np.arange(256)).all()


def test_text_size_against_the_golden_advances(atlas):
    gold = np.load(os.path.join(HERE, "golden", "hershey_atlas_cv2.npz"))
    adv = gold["advances"]
    for text in ("", "A", "person: 0.87", "traffic light: 0.99", "~{|}", "\x01\xff"):
        units = sum(int(adv[(b if 32 <= b <= 126 else ord("?")) - 32]) for b in text.encode("latin1"))
        assert atlas.text_size(text.encode("latin1").decode("latin1")) is not None
        if all(ord(c) < 128 for c in text):
            assert atlas.text_size(text) == (int(round(units * 0.75 + 2)), int(gold["text_height"]))


def _label(text, x, y):
    """An overlay whose label origin lands exactly at (x, y): a name-less rectangle cannot carry text,
    so the op list is taken from a label placed by hand through the rule's third branch."""
    return {"objects": [{"name": text, "confidence": 0.5}], "rectangles": [{"x": x, "y": y, "w": 0, "h": 0}]}


def test_random_labels_vs_cv2_putText_everywhere(atlas):
    """10 000 random label placements, most of them cut by an image border or a corner: replaying
    the planner's glyph ops must equal cv2.putText bit for bit (cv2 is the oracle, never the product)."""
    cv2 = pytest.importorskip("cv2")
    from aiko_services_b200.overlay_plan import plan
    rng = np.random.default_rng(11)
    chars = [chr(c) for c in range(32, 127)]
    names = ["person", "dog", "traffic light", "Zq{|}~", "W@#", "i.l'"]
    done = 0
    while done < 10000:
        h, w = int(rng.integers(8, 120)), int(rng.integers(8, 260))
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        name = names[int(rng.integers(len(names)))] if rng.random() < 0.5 else \
            "".join(rng.choice(chars, int(rng.integers(1, 8))))
        name = name.strip() or "x"
        conf = float(rng.integers(1, 100)) / 100
        # boxes hugging every border so that labels get clipped left / right / top / bottom / corners
        x = int(rng.integers(-30, w + 5)); y = int(rng.integers(-10, h + 10))
        bw, bh = int(rng.integers(0, 40)), int(rng.integers(0, 30))
        overlay = {"objects": [{"name": name, "confidence": conf}], "rectangles": [{"x": x, "y": y, "w": bw, "h": bh}]}
        text = f"{name}: {conf:0.2f}"
        (tw, th), _ = cv2.getTextSize(text, cv2.FONT_HERSHEY_SIMPLEX, 0.75, 2)
        tx, ty = x, (int(y - th / 2 - 2) if y > th * 1.5 else y + bh + th + 2)
        if x + tw > w:
            tx, ty = x - tw, y + th
        ref = img.copy()
        cv2.putText(ref, text, (tx, ty), cv2.FONT_HERSHEY_SIMPLEX, 0.75, (255, 255, 0), 2, cv2.LINE_AA)
        cv2.rectangle(ref, (x, y), (x + bw, y + bh), (255, 255, 0), 2)
        got = replay_ops_cpu(img.copy(), plan(overlay, w, h, "rgb").ops()[0], atlas)
        np.testing.assert_array_equal(got, ref, err_msg=f"{text!r} at ({tx},{ty}) in {w}x{h}")
        done += 1


def test_no_opencv_in_the_product():
    import re
    root = os.path.join(HERE, "..", "aiko_services_b200")
    hits = []
    for d, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(d, f)).read()
                for m in re.finditer(r"^\s*(import cv2|from cv2)", text, re.M):
                    hits.append((f, text[:m.start()].count("\n") + 1))
    # cv2 may only appear in elements that wrap OpenCV's *codecs / capture* (video_io, webcam_io)
    assert all(f in ("video_io.py", "webcam_io.py") for f, _ in hits), hits
