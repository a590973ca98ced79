"""ctypes wrapper over oracle/liboracle.so (see image_oracle.c for citations).

TEST INFRASTRUCTURE.  Every function takes / returns numpy arrays shaped like the
reference elements' images: uint8 [H, W, 3] or [H, W].
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    sources = [os.path.join(_HERE, name) for name in ("image_oracle.c", "jpeg_oracle.c")]
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(s) for s in sources):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _hwc(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim == 2:
        return img, img.shape[0], img.shape[1], 1
    return img, img.shape[0], img.shape[1], img.shape[2]


def resize_linear(img, width, height):
    """cv2.resize(img, (width, height)) -- R/elements/media/image_io.py:216"""
    img, h, w, c = _hwc(img)
    out = np.empty((height, width) + img.shape[2:], np.uint8)
    rc = lib().orc_resize_linear_u8(_p(img), h, w, c, _p(out), height, width)
    assert rc == 0
    return out


def resize_bicubic_aa(img, width, height):
    """PIL.Image.resize((width, height)) -- R/elements/media/image_io.py:218"""
    img, h, w, c = _hwc(img)
    out = np.empty((height, width) + img.shape[2:], np.uint8)
    rc = lib().orc_resize_bicubic_aa_u8(_p(img), h, w, c, _p(out), height, width)
    assert rc == 0
    return out


def swap_rb(img):
    img, h, w, c = _hwc(img)
    assert c == 3
    out = np.empty_like(img)
    lib().orc_swap_rb_u8(_p(img), _p(out), ctypes.c_size_t(h * w))
    return out


def bgr2gray(img, rgb_order=False):
    img, h, w, c = _hwc(img)
    assert c == 3
    out = np.empty((h, w), np.uint8)
    lib().orc_bgr2gray_u8(_p(img), _p(out), ctypes.c_size_t(h * w), int(rgb_order))
    return out


def gray2bgr(img):
    img, h, w, c = _hwc(img)
    assert c == 1
    out = np.empty((h, w, 3), np.uint8)
    lib().orc_gray2bgr_u8(_p(img), _p(out), ctypes.c_size_t(h * w))
    return out


def flip(img, flip_code):
    img, h, w, c = _hwc(img)
    out = np.empty_like(img)
    lib().orc_flip_u8(_p(img), _p(out), h, w, c, int(flip_code))
    return out


def rectangle(img, x0, y0, x1, y1, colour):
    """In place, like cv2.rectangle(img, (x0,y0), (x1,y1), colour, 2)."""
    assert img.dtype == np.uint8 and img.flags.c_contiguous
    h, w = img.shape[:2]
    c = 1 if img.ndim == 2 else img.shape[2]
    col = np.asarray(colour, np.uint8)[:c].copy()
    lib().orc_rectangle_u8(_p(img), h, w, c, int(x0), int(y0), int(x1), int(y1), _p(col))
    return img


def text_lut(img, px, py, patches, luts):
    """Stamp per-channel transfer-function patches (see text atlas) in place."""
    assert img.dtype == np.uint8 and img.flags.c_contiguous
    h, w = img.shape[:2]
    c = 1 if img.ndim == 2 else img.shape[2]
    ph, pw = patches[0].shape
    pp = (ctypes.c_void_p * c)(*[_p(np.ascontiguousarray(p, np.uint16)) for p in patches])
    keep = [np.ascontiguousarray(p, np.uint16) for p in patches]
    pp = (ctypes.c_void_p * c)(*[_p(k) for k in keep])
    lk = [np.ascontiguousarray(l, np.uint8) for l in luts]
    lp = (ctypes.c_void_p * c)(*[_p(k) for k in lk])
    lib().orc_text_lut_u8(_p(img), h, w, c, int(px), int(py), ph, pw, pp, lp)
    return img


def resize_norm_f32(img, width, height, mean, std):
    img, h, w, c = _hwc(img)
    assert c == 3
    out = np.empty((3, height, width), np.float32)
    m = np.asarray(mean, np.float32)
    s = np.asarray(std, np.float32)
    rc = lib().orc_resize_norm_f32(_p(img), h, w, _p(out), height, width, _p(m), _p(s))
    assert rc == 0
    return out


# ---- PIL.Image.resize for modes other than RGB / L (numpy restatement of Pillow's Convert.c / Geometry.c)
def _muldiv255(a, b):
    t = a.astype(np.uint32) * b.astype(np.uint32) + 128
    return (((t >> 8) + t) >> 8).astype(np.uint8)


def pil_premultiply(img):
    """RGBA -> RGBa / LA -> La (Pillow rgbA2rgba, lA2la); LA comes back as 4 bands (L, L, L, A)."""
    img = np.asarray(img, np.uint8)
    if img.shape[2] == 2:
        img = np.stack([img[:, :, 0]] * 3 + [img[:, :, 1]], axis=2)
    out = img.copy()
    for c in range(3):
        out[:, :, c] = _muldiv255(img[:, :, c], img[:, :, 3])
    return out


def pil_unpremultiply(img, bands):
    """RGBa -> RGBA / La -> LA (Pillow rgba2rgbA, la2lA): CLIP8(255 * v / alpha) unless alpha is 0 or 255."""
    img = np.asarray(img, np.uint8)
    a = img[:, :, 3].astype(np.uint32)
    out = img.copy()
    keep = (a == 0) | (a == 255)
    for c in range(3):
        q = np.minimum((255 * img[:, :, c].astype(np.uint32)) // np.maximum(a, 1), 255)
        out[:, :, c] = np.where(keep, img[:, :, c], q).astype(np.uint8)
    return out if bands == 4 else np.stack([out[:, :, 0], out[:, :, 3]], axis=2)


def pil_nearest_table(src, dst):
    """ImagingScaleAffine: position 0.5 * a, then += a per output index (a = src / dst, double)."""
    a = src / dst
    pos, tab = a * 0.5, []
    for _ in range(dst):
        i = -1 if pos < 0.0 else int(pos)
        tab.append(i if 0 <= i < src else -1)
        pos += a
    return np.array(tab)


def pil_resize(img, mode, width, height):
    """PIL.Image.resize((width, height)) on the array of an image of PIL mode ``mode``
    (R/elements/media/image_io.py:218 for whatever mode the file had)."""
    img = np.asarray(img, np.uint8)
    if mode in ("LA", "RGBA"):
        return pil_unpremultiply(resize_bicubic_aa(pil_premultiply(img), width, height), img.shape[2])
    if mode in ("P", "1"):
        xt, yt = pil_nearest_table(img.shape[1], width), pil_nearest_table(img.shape[0], height)
        out = img[np.clip(yt, 0, None)][:, np.clip(xt, 0, None)].copy()
        out[yt < 0] = 0
        out[:, xt < 0] = 0
        return out
    return resize_bicubic_aa(img, width, height)


# ---- RGB / BGR <-> YUV (extension: no reference call site; numpy restatement of cv2.cvtColor's 8-bit
# arithmetic, pinned against cv2 over all 2^24 colours in tests/test_oracle_golden.py)
def _descale14(x):
    return (x + (1 << 13)) >> 14


def yuv(img, bgr_order=False, inverse=False):
    a = np.asarray(img, np.uint8).astype(np.int64)
    if not inverse:
        r, g, b = (a[..., 2], a[..., 1], a[..., 0]) if bgr_order else (a[..., 0], a[..., 1], a[..., 2])
        y = _descale14(r * 4899 + g * 9617 + b * 1868)
        u = _descale14((b - y) * 8061 + (128 << 14))
        v = _descale14((r - y) * 14369 + (128 << 14))
        out = np.stack([y, u, v], -1)
    else:
        y, u, v = a[..., 0], a[..., 1] - 128, a[..., 2] - 128
        r = y + _descale14(v * 18678)
        g = y + _descale14(v * -9519 + u * -6472)
        b = y + _descale14(u * 33292)
        out = np.stack([b, g, r] if bgr_order else [r, g, b], -1)
    return np.clip(out, 0, 255).astype(np.uint8)


def jpeg_decode(data):
    """Baseline JPEG -> uint8 [H, W, 3] RGB, as libjpeg-turbo's default decoder (what Pillow runs) does;
    gray files give R = G = B.  Raises ValueError for files outside the oracle's subset."""
    data = bytes(data)
    w, h = ctypes.c_int(), ctypes.c_int()
    rc = lib().orc_jpeg_decode_rgb(data, len(data), None, ctypes.byref(w), ctypes.byref(h))
    if rc != 0:
        raise ValueError(f"jpeg oracle: cannot read this file (code {rc})")
    out = np.empty((h.value, w.value, 3), np.uint8)
    rc = lib().orc_jpeg_decode_rgb(data, len(data), _p(out), ctypes.byref(w), ctypes.byref(h))
    if rc != 0:
        raise ValueError(f"jpeg oracle: cannot decode this file (code {rc})")
    return out


def jpeg_encode(img, quality=75, subsampling=(2, 2)):
    """uint8 [H, W, 3] RGB or [H, W] gray -> the bytes PIL.Image.fromarray(img).save(f, "JPEG", quality=...,
    subsampling=...) writes (libjpeg-turbo defaults: standard Huffman tables, JFIF header).
    ``subsampling``: luma sampling factors (2, 2) = 4:2:0 (Pillow's default), (2, 1) = 4:2:2, (1, 1) = 4:4:4."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape[:2]
    channels = 1 if img.ndim == 2 else img.shape[2]
    fn = lib().oracle_jpeg_encode
    fn.restype = ctypes.c_size_t
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                   ctypes.c_void_p, ctypes.c_size_t]
    cap = h * w * channels * 2 + 4096
    out = np.empty(cap, np.uint8)
    n = fn(img.ctypes.data, h, w, channels, int(quality), int(subsampling[0]), int(subsampling[1]), out.ctypes.data, cap)
    if n > cap:
        out = np.empty(n, np.uint8)
        n = fn(img.ctypes.data, h, w, channels, int(quality), int(subsampling[0]), int(subsampling[1]), out.ctypes.data, n)
    return out[:n].tobytes()


def pil_paste(img, rgba, x, y):
    """PIL's ``Image.paste(image, (x, y), image)`` of an RGBA image onto an RGB frame (Paste.c BLEND8): per byte
    t = dst * (255 - a) + src * a + 128, out = ((t >> 8) + t) >> 8; the box is clipped to the frame."""
    out = np.array(img, dtype=np.uint8, copy=True)
    h, w = out.shape[:2]
    sh, sw = rgba.shape[:2]
    x0, y0, x1, y1 = max(x, 0), max(y, 0), min(x + sw, w), min(y + sh, h)
    if x1 <= x0 or y1 <= y0:
        return out
    part = rgba[y0 - y:y1 - y, x0 - x:x1 - x].astype(np.int32)
    a = part[..., 3:4]
    t = out[y0:y1, x0:x1].astype(np.int32) * (255 - a) + part[..., :3] * a + 128
    out[y0:y1, x0:x1] = (((t >> 8) + t) >> 8).astype(np.uint8)
    return out
