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
    src = os.path.join(_HERE, "image_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
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
