"""Seeded inputs shared by make_golden.py (which runs the reference) and the parity tests."""
import numpy as np

OVERLAY_4 = {   # 4 rectangles + 4 labels: one clipped by the border, one flipped by the x+tw rule
    "objects": [{"name": "person", "confidence": 0.87}, {"name": "bicycle", "confidence": 0.5},
                {"name": "dog", "confidence": 0.33}, {"name": "traffic light", "confidence": 0.99}],
    "rectangles": [{"x": 100, "y": 120, "w": 200, "h": 150}, {"x": 300, "y": 10, "w": 120, "h": 90},
                   {"x": 560, "y": 300, "w": 120, "h": 100}, {"x": 20, "y": 400, "w": 260, "h": 120}],
}
OVERLAY_SMALL = {
    "objects": [{"name": "cat", "confidence": 0.61}, {"name": "tv", "confidence": 0.0},
                {"confidence": 0.9}, {"name": "Zq", "confidence": 0.05}],
    "rectangles": [{"x": 40, "y": 50, "w": 60, "h": 40}, {"x": 5, "y": 5, "w": 20, "h": 20},
                   {"x": 120, "y": 20, "w": 50, "h": 70}, {"x": 150, "y": 100, "w": 30, "h": 15}],
}
OVERLAY_RECTS_ONLY = {"rectangles": [{"x": -5, "y": -5, "w": 30, "h": 30}, {"x": 60, "y": 40, "w": 0, "h": 0},
                                     {"x": 100, "y": 60, "w": 300, "h": 300}, {"x": 30, "y": 70, "w": 2, "h": 3}]}

# (name, seed, src_h, src_w, channels, dst_w, dst_h)
RESIZE_NDARRAY = [
    ("down_3x", 11, 108, 192, 3, 64, 48),
    ("down_odd", 12, 97, 131, 3, 50, 41),
    ("up", 13, 30, 40, 3, 100, 70),
    ("area2x", 14, 64, 96, 3, 48, 32),
    ("gray", 15, 75, 99, 1, 33, 60),
    ("one_pixel", 16, 1, 1, 3, 7, 5),
    ("to_one", 17, 40, 50, 3, 1, 1),
    ("rgba", 18, 45, 61, 4, 90, 20),
    ("cfg2_1080p", 2000, 1080, 1920, 3, 640, 480),
    ("cfg4_4k_224", 4000, 2160, 3840, 3, 224, 224),
]
RESIZE_PIL = [
    ("pil_down", 21, 108, 192, 64, 48),
    ("pil_up", 22, 30, 40, 100, 70),
    ("pil_mixed", 23, 90, 50, 33, 120),
    ("pil_same_w", 24, 60, 64, 64, 20),
    ("cfg1_1080p", 1000, 1080, 1920, 640, 480),
]
# (name, seed, h, w, channels, overlay)
OVERLAY = [
    ("ov4_640x480", 31, 480, 640, 3, OVERLAY_4),
    ("ov_small", 32, 128, 192, 3, OVERLAY_SMALL),
    ("ov_gray", 33, 128, 192, 1, OVERLAY_SMALL),
    ("ov_rects", 34, 128, 192, 3, OVERLAY_RECTS_ONLY),
]


def frame(seed, h, w, c):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
    return img[:, :, 0].copy() if c == 1 else img

# PIL modes other than RGB through ImageResize's PIL branch: (name, seed, mode, src_h, src_w, dst_w, dst_h)
RESIZE_PIL_MODES = [
    ("rgba_down", 31, "RGBA", 108, 192, 64, 48),
    ("rgba_up", 32, "RGBA", 30, 41, 100, 70),
    ("la_down", 33, "LA", 90, 131, 50, 41),
    ("la_up", 34, "LA", 20, 30, 64, 48),
    ("l_down", 35, "L", 97, 131, 50, 41),
    ("cmyk_down", 36, "CMYK", 64, 96, 40, 30),
    ("p_down", 37, "P", 108, 192, 64, 48),
    ("p_up_odd", 38, "P", 31, 47, 101, 67),
    ("bilevel_down", 39, "1", 99, 150, 60, 40),
    ("bilevel_up", 40, "1", 20, 33, 77, 50),
]


def pil_image(seed, mode, h, w):
    """A seeded PIL image of ``mode``; alpha channels mix 0, 255 and everything between."""
    from PIL import Image
    rng = np.random.default_rng(seed)
    bands = {"L": 1, "P": 1, "1": 1, "LA": 2, "RGB": 3, "RGBA": 4, "CMYK": 4}[mode]
    a = rng.integers(0, 256, (h, w, bands), dtype=np.uint8)
    if mode in ("LA", "RGBA"):
        alpha = a[:, :, -1]
        alpha[rng.random((h, w)) < 0.2] = 0
        alpha[rng.random((h, w)) < 0.2] = 255
    if mode == "1":
        return Image.fromarray(a[:, :, 0] > 127)
    if bands == 1:
        im = Image.fromarray(a[:, :, 0], "L" if mode == "L" else "P")
        if mode == "P":
            im.putpalette(rng.integers(0, 256, 768, dtype=np.uint8).tobytes())
        return im
    return Image.fromarray(a, mode)
