"""CPU: the draw-list builder + glyph atlas, replayed with the C oracle, against the golden
ImageOverlay outputs of the reference and against cv2 itself on random overlays."""
import numpy as np
import pytest

from aiko_services_b200.overlay_plan import build_ops, label_origin
from aiko_services_b200.text_atlas import TextAtlas
from golden import cases
from oracle import reference_port
from util import check_golden, overlay_cpu, random_overlay


@pytest.fixture(scope="module")
def atlas():
    return TextAtlas()


@pytest.mark.parametrize("case", cases.OVERLAY, ids=lambda c: c[0])
def test_overlay_golden(golden, atlas, case):
    name, seed, h, w, c, overlay = case
    check_golden(golden, "overlay/" + name, overlay_cpu(cases.frame(seed, h, w, c), overlay, atlas))


def test_text_size_matches_cv2(atlas):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    chars = [chr(c) for c in range(32, 127)]
    for _ in range(300):
        text = "".join(rng.choice(chars, int(rng.integers(0, 30))))
        (tw, th), _ = cv2.getTextSize(text, cv2.FONT_HERSHEY_SIMPLEX, 0.75, 2)
        assert atlas.text_size(text) == (tw, th), text


def test_interior_labels_bit_exact_vs_cv2(atlas):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(4)
    chars = [chr(c) for c in range(32, 127)]
    for _ in range(60):
        h, w = int(rng.integers(80, 160)), int(rng.integers(420, 600))
        img = cases.frame(int(rng.integers(1 << 30)), h, w, 3)
        text = "".join(rng.choice(chars, int(rng.integers(1, 16))))
        overlay = {"objects": [{"name": text, "confidence": float(rng.random() * 0.98 + 0.01)}],
                   "rectangles": [{"x": int(rng.integers(10, 60)), "y": int(rng.integers(40, h - 30)),
                                   "w": 40, "h": 20}]}
        ref = reference_port.image_overlay([img], overlay)[0]
        np.testing.assert_array_equal(overlay_cpu(img, overlay, atlas), ref, err_msg=repr(text))


def test_random_overlays_incl_border_clipping_vs_cv2(atlas):
    pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    for i in range(60):
        h, w = int(rng.integers(40, 200)), int(rng.integers(60, 320))
        c = 1 if i % 5 == 0 else 3
        img = cases.frame(int(rng.integers(1 << 30)), h, w, c)
        overlay = random_overlay(rng, w, h)
        ref = reference_port.image_overlay([img], overlay)[0]
        np.testing.assert_array_equal(overlay_cpu(img, overlay, atlas), ref, err_msg=str(overlay))


def test_placement_rule(atlas):
    # label above the box, below it when near the top, flipped left when it would overflow
    assert label_origin(atlas, "a: 0.50", 100, 120, 50, 40, 640) == (100, int(120 - 17 / 2 - 2))
    assert label_origin(atlas, "a: 0.50", 100, 10, 50, 40, 640) == (100, 10 + 40 + 17 + 2)
    tw, th = atlas.text_size("a: 0.50")
    assert label_origin(atlas, "a: 0.50", 600, 120, 50, 40, 640) == (600 - tw, 120 + th)


def test_skips_and_empty(atlas):
    assert build_ops({}, 64, 48) == []
    assert build_ops({"rectangles": []}, 64, 48) == []
    ops = build_ops({"objects": [{"name": "a", "confidence": 0.0}], "rectangles": [{"x": 1, "y": 1, "w": 5, "h": 5}]},
                    64, 48)
    assert ops == []      # confidence must exceed the threshold (image_io.py:131)


def test_marks_cover_every_pixel_an_op_changes(atlas):
    """The GPU applies an op only inside its marks (blocks of 4-pixel groups, csrc/overlay_marks.h):
    for random overlays (clipped boxes, labels at the border), every pixel an op changes when
    replayed alone must lie in one of its marks, and the planner's mark / cell bounds must hold."""
    import ctypes
    from aiko_services_b200 import _native
    from aiko_services_b200.overlay_plan import plan
    from util import replay_ops_cpu
    lib = _native.lib()
    rng = np.random.default_rng(6)
    for trial in range(12):
        h, w = int(rng.integers(40, 140)), int(rng.integers(60, 260))
        overlay = random_overlay(rng, w, h, n=int(rng.integers(1, 7)))
        planned = plan(overlay, w, h, "rgb")
        ops = planned.ops()[0]
        n_marks = n_cells = 0
        for op in ops:
            marks = (ctypes.c_int16 * 16)()
            n = lib.aiko_b200_overlay_marks_host(ctypes.byref(op), w, h, marks)
            assert 0 <= n <= 4
            blocks = [tuple(marks[4 * i:4 * i + 4]) for i in range(n)]
            n_marks += n
            n_cells += sum((y1 - y0 + 1) * (g1 - g0 + 1) for y0, y1, g0, g1 in blocks)
            for y0, y1, g0, g1 in blocks:
                assert 0 <= y0 <= y1 < h and 0 <= g0 <= g1 <= (w - 1) // 4
            probes = [np.zeros((h, w, 3), np.uint8), np.full((h, w, 3), 255, np.uint8),
                      rng.integers(0, 256, (h, w, 3), dtype=np.uint8)]
            changed = np.zeros((h, w), bool)
            for base in probes:
                changed |= (replay_ops_cpu(base.copy(), [op], atlas) != base).any(axis=2)
            for y, x in zip(*np.nonzero(changed)):
                assert any(y0 <= y <= y1 and g0 <= x // 4 <= g1 for y0, y1, g0, g1 in blocks), (trial, int(x), int(y))
        assert planned.info.max_marks_per_list == n_marks and planned.info.max_cells_per_list == n_cells
        assert planned.info.n_ops == planned.info.max_ops_per_list == len(ops)


def test_shared_and_per_frame_lists(atlas):
    from aiko_services_b200.overlay_plan import plan
    a, b = cases.OVERLAY_4, {"rectangles": [{"x": 3, "y": 4, "w": 10, "h": 12}]}
    p = plan([a, b, a], 640, 480, "bgr")
    assert p.n_lists == 2 and p.n_frames == 3
    tail = p.buffer[p.ops_bytes:].view(np.int32)
    assert list(tail[p.n_lists + 1:]) == [0, 1, 0]
    assert tail[0] == 0 and tail[2] == p.n_ops and len(p.ops()[1]) == 1
