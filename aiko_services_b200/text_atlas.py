"""Glyph atlas for ImageOverlay's anti-aliased labels.

The reference draws labels with ``cv2.putText(image, text, (tx, ty), FONT_HERSHEY_SIMPLEX,
0.75, (0, 255, 255), 2, cv2.LINE_AA)`` (R/elements/media/image_io.py:101-105,151-153).  OpenCV
rasterises one character at a time; every stroke pixel is a read-modify-write that depends
only on the pixel's current value and the colour component, and the pen advances in
multiples of 0.75 px, so a character only ever lands on four sub-pixel phases.  The atlas
therefore stores, per (character, phase, colour component), a small patch of per-pixel
transfer-function ids into a table of 256-entry look-up tables; replaying the characters of
a label in order through those tables reproduces cv2.putText bit for bit (tests/
test_text_atlas.py), with no OpenCV call on the per-frame path.

The font outlines are OpenCV's Hershey data, which the reference does not vendor: the
atlas file ``data/hershey_simplex_s075_t2.npz`` is generated once by ``bake()`` below (needs
cv2) and committed.  cv2 clips anti-aliased strokes against the image rectangle, which
changes coverage within ``BORDER_MARGIN`` pixels of the border; glyphs that close to the
border are baked on demand against the actual border geometry (``border_variant``).
"""
import os

import numpy as np

FONT_SCALE = 0.75   # image_io.py:103
THICKNESS = 2       # image_io.py:104
COMPONENTS = (0, 255)   # colour components of (0, 255, 255), image_io.py:101
FIRST_CHAR, LAST_CHAR = 32, 126
PAD_X = 8           # patch origin = (floor(pen_x) - PAD_X, baseline_y - PAD_Y)
PAD_Y = 40
PATCH_W, PATCH_H = 40, 48
BORDER_MARGIN = 6   # strokes nearer than this to the image border may be clipped by cv2

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "hershey_simplex_s075_t2.npz")


def _phase_prefixes(cv2, advances):
    """Strings whose total advance puts the pen on each quarter-pixel phase, followed by
    two spaces (24 px) so their strokes stay clear of the baked character."""
    chars = [chr(c) for c in range(FIRST_CHAR + 1, LAST_CHAR + 1)]
    prefixes = {0: ""}
    for units_mod4 in (1, 2, 3):
        found = None
        for n in (1, 2, 3):
            for ch in chars:
                if (advances[ord(ch) - FIRST_CHAR] * n) % 4 == units_mod4:
                    found = ch * n
                    break
            if found:
                break
        assert found, "no prefix reaches this pen phase"
        prefixes[(3 * units_mod4) % 4] = found   # pen = 0.75 * units -> quarter pixels = 3 * units
    return {q: (p + "  " if p else "") for q, p in prefixes.items()}


def _render_transfer(cv2, text, org, width, height, colour_bgr):
    """-> uint8[256, height, width, 3]: text rendered on every constant background."""
    out = np.empty((256, height, width, 3), np.uint8)
    for b in range(256):
        img = np.full((height, width, 3), b, np.uint8)
        cv2.putText(img, text, org, cv2.FONT_HERSHEY_SIMPLEX, FONT_SCALE, colour_bgr, THICKNESS, cv2.LINE_AA)
        out[b] = img
    return out


def bake(path=_DATA):
    """Generate the interior atlas with OpenCV.  Run once; the result is committed."""
    import cv2
    font = cv2.FONT_HERSHEY_SIMPLEX
    n_chars = LAST_CHAR - FIRST_CHAR + 1
    advances = np.array([cv2.getTextSize(chr(FIRST_CHAR + i), font, 1.0, 0)[0][0] for i in range(n_chars)],
                        np.int32)
    text_height = cv2.getTextSize("A", font, FONT_SCALE, THICKNESS)[0][1]
    prefixes = _phase_prefixes(cv2, advances)
    identity = np.arange(256, dtype=np.uint8)
    lut_index = {identity.tobytes(): 0}
    luts = [identity]
    n_patches = 1 + n_chars * 4 * len(COMPONENTS)
    patch_ids = np.zeros((n_patches, PATCH_H, PATCH_W), np.uint16)
    bbox = np.zeros((n_patches, 4), np.int16)
    bbox[:, 0] = bbox[:, 1] = 1   # empty: x1 < x0
    org_x, org_y = 16, PAD_Y
    for ci in range(n_chars):
        for q in range(4):
            prefix = prefixes[q]
            units = sum(int(advances[ord(c) - FIRST_CHAR]) for c in prefix)
            pen_q = 3 * units
            assert pen_q % 4 == q
            x0 = org_x + pen_q // 4 - PAD_X
            width = x0 + PATCH_W + 16
            # colour (0, 255, 255): channel 0 gives component 0, channel 1 gives component 255
            tr = _render_transfer(cv2, prefix + chr(FIRST_CHAR + ci), (org_x, org_y), width, PATCH_H + 8,
                                  (0, 255, 255))
            region = tr[:, :PATCH_H, x0:x0 + PATCH_W, :]
            # nothing of this character may fall outside the patch window
            outside = tr.copy()
            outside[:, :PATCH_H, x0:x0 + PATCH_W, :] = identity[:, None, None, None]
            if prefix:
                outside[:, :, :x0 - 8, :] = identity[:, None, None, None]
            assert (outside == identity[:, None, None, None]).all(), f"glyph {chr(FIRST_CHAR + ci)!r} leaves its patch"
            for k, comp in enumerate(COMPONENTS):
                chan = 0 if comp == 0 else 1
                pidx = 1 + (ci * 4 + q) * len(COMPONENTS) + k
                fn = region[:, :, :, chan]                    # [256, PH, PW]
                touched = (fn != identity[:, None, None]).any(0)
                ys, xs = np.nonzero(touched)
                if len(ys):
                    bbox[pidx] = (xs.min(), ys.min(), xs.max(), ys.max())
                for y, x in zip(ys, xs):
                    key = fn[:, y, x].tobytes()
                    idx = lut_index.get(key)
                    if idx is None:
                        idx = len(luts)
                        lut_index[key] = idx
                        luts.append(fn[:, y, x].copy())
                    patch_ids[pidx, y, x] = idx
    luts = np.stack(luts).astype(np.uint8)
    assert len(luts) < 65536
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, advances=advances, text_height=np.int32(text_height), patch_ids=patch_ids,
                        bbox=bbox, luts=luts, cv2_version=np.array(cv2.__version__))
    return path


class TextAtlas:
    """Host copy of the atlas plus label layout; ``device(dev)`` gives the GPU-resident tables."""

    _shared = None

    def __init__(self, path=_DATA):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run aiko_services_b200.text_atlas.bake() where cv2 is installed")
        z = np.load(path)
        self.advances = z["advances"].astype(np.int64)
        self.text_height = int(z["text_height"])
        self.patch_ids = z["patch_ids"]
        self.bbox = z["bbox"].astype(np.int64)
        self.luts = z["luts"]
        self.cv2_version = str(z["cv2_version"])
        self.patch_w, self.patch_h = PATCH_W, PATCH_H
        self._lut_index = None          # lazily: bytes -> id, for border variants
        self._variants = {}             # (char, phase, comp, border key) -> patch index
        self._device = {}               # torch.device -> _DeviceAtlas
        self.version = 0                # bumped whenever patches / luts grow

    @classmethod
    def shared(cls):
        if cls._shared is None:
            cls._shared = cls()
        return cls._shared

    # -- layout (cv2.getTextSize + the reference's placement rule) -------------------
    def text_size(self, text):
        """cv2.getTextSize(text, FONT_HERSHEY_SIMPLEX, 0.75, 2)[0]"""
        units = sum(int(self.advances[self._ci(ch)]) for ch in text)
        return int(round(units * FONT_SCALE + THICKNESS)), self.text_height   # round(): half to even, as cvRound

    @staticmethod
    def _ci(ch):
        c = ord(ch)
        if c < FIRST_CHAR or c > LAST_CHAR:
            c = ord("?")       # OpenCV draws '?' for characters outside the font
        return c - FIRST_CHAR

    def patch_index(self, ch, phase, comp):
        return 1 + (self._ci(ch) * 4 + phase) * len(COMPONENTS) + COMPONENTS.index(comp)

    def glyphs(self, text, org_x, org_y):
        """Yield (char, phase, patch_origin_x, patch_origin_y) for cv2.putText(text, (org_x, org_y))."""
        units = 0
        for ch in text:
            pen_q = 3 * units
            yield ch, pen_q % 4, org_x + pen_q // 4 - PAD_X, org_y - PAD_Y
            units += int(self.advances[self._ci(ch)])

    # -- border variants ------------------------------------------------------------------
    def border_key(self, pidx, px, py, width, height):
        """None if the glyph stamped at (px, py) stays clear of the border band, else a key
        describing where the image borders lie relative to the patch."""
        x0, y0, x1, y1 = self.bbox[pidx]
        if x1 < x0:
            return None
        m = BORDER_MARGIN
        if px + x0 >= m and py + y0 >= m and px + x1 < width - m and py + y1 < height - m:
            return None
        far = 1 << 20
        left = px if px + x0 < m else -far
        top = py if py + y0 < m else -far
        right = width - px if px + x1 >= width - m else far
        bottom = height - py if py + y1 >= height - m else far
        return (left, top, right, bottom)

    def border_variant(self, ch, phase, comps, key):
        """Patch indices (one per requested component) for a glyph baked against image borders
        at the offsets in ``key``.  Needs cv2 (font outlines); results are cached."""
        missing = [c for c in comps if (ch, phase, c, key) not in self._variants]
        if missing:
            self._bake_variant(ch, phase, key)
        return [self._variants[(ch, phase, c, key)] for c in comps]

    def _bake_variant(self, ch, phase, key):
        try:
            import cv2
        except ModuleNotFoundError as e:   # loud: there is no silent approximation
            raise RuntimeError("a label touches the image border and its clipped glyphs must be baked with "
                               "OpenCV's font data, but cv2 is not installed") from e
        if self._lut_index is None:
            self._lut_index = {self.luts[i].tobytes(): i for i in range(len(self.luts))}
        left, top, right, bottom = key
        far = 1 << 19
        slack = PATCH_W + 64
        # image rectangle in patch coordinates: [-left', right') x [-top', bottom')
        img_x0 = -left if left > -far else -slack     # patch x of image column 0 ... px = left
        img_y0 = -top if top > -far else -slack
        img_x1 = right if right < far else PATCH_W + slack
        img_y1 = bottom if bottom < far else PATCH_H + slack
        width, height = img_x1 - img_x0, img_y1 - img_y0
        identity = np.arange(256, dtype=np.uint8)
        new_ids = {c: np.zeros((PATCH_H, PATCH_W), np.uint16) for c in COMPONENTS}
        if width > 0 and height > 0:
            prefixes = _phase_prefixes(cv2, self.advances)
            prefix = prefixes[phase]
            units = sum(int(self.advances[self._ci(c)]) for c in prefix)
            pen_q = 3 * units
            # patch origin in image coords is (-img_x0, -img_y0); pen floor = origin + PAD_X
            org_x = -img_x0 + PAD_X - pen_q // 4
            org_y = -img_y0 + PAD_Y
            tr = _render_transfer(cv2, prefix + ch, (org_x, org_y), width, height, (0, 255, 255))
            for comp in COMPONENTS:
                chan = 0 if comp == 0 else 1
                ids = new_ids[comp]
                for y in range(max(0, img_y0), min(PATCH_H, img_y1)):
                    for x in range(max(0, img_x0), min(PATCH_W, img_x1)):
                        fn = tr[:, y - img_y0, x - img_x0, chan]
                        if (fn != identity).any():
                            k = fn.tobytes()
                            idx = self._lut_index.get(k)
                            if idx is None:
                                idx = len(self.luts)
                                self._lut_index[k] = idx
                                self.luts = np.concatenate([self.luts, fn[None].copy()])
                            ids[y, x] = idx
        assert len(self.luts) < 65536
        for comp in COMPONENTS:
            pidx = len(self.patch_ids)
            self.patch_ids = np.concatenate([self.patch_ids, new_ids[comp][None]])
            ys, xs = np.nonzero(new_ids[comp])
            box = (xs.min(), ys.min(), xs.max(), ys.max()) if len(ys) else (1, 1, 0, 0)
            self.bbox = np.concatenate([self.bbox, np.array([box], np.int64)])
            self._variants[(ch, phase, comp, key)] = pidx
        self.version += 1

    # -- device copy ---------------------------------------------------------------------------
    def device(self, device):
        import torch
        device = torch.device(device)
        d = self._device.get(device)
        if d is None or d.version != self.version:
            d = _DeviceAtlas(self, device)
            self._device[device] = d
        return d


class _DeviceAtlas:
    def __init__(self, atlas, device):
        import torch
        self.version = atlas.version
        self.patch_w, self.patch_h = atlas.patch_w, atlas.patch_h
        self.patch_ids = torch.from_numpy(np.ascontiguousarray(atlas.patch_ids.view(np.int16))).to(device)
        self.luts = torch.from_numpy(np.ascontiguousarray(atlas.luts)).to(device)
