"""Python-facing operators over the C ABI.

torch is used here for device memory and streams only (``tensor.data_ptr()``,
``torch.cuda.current_stream()``); the arithmetic is in libaiko_b200.so.  Each function
mirrors the library call the reference element makes (cited in include/aiko_b200.h) but
takes a whole batch ``uint8[N, H, W, C]`` (or ``[N, H, W]`` for grayscale) resident on the GPU.
"""
import ctypes
import os

import numpy as np
import torch

from . import _native
from ._native import Overlay, OverlayOp, check


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


def _require_cuda(t, name, dtype=torch.uint8):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor (aiko_services_b200 has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: expected a contiguous tensor")
    return t


def _nhwc(t):
    """-> (n, h, w, c, gray_input)"""
    if t.dim() == 4:
        return t.shape[0], t.shape[1], t.shape[2], t.shape[3], False
    if t.dim() == 3:
        return t.shape[0], t.shape[1], t.shape[2], 1, True
    raise ValueError(f"expected [N,H,W,C] or [N,H,W], got {tuple(t.shape)}")


def _out(out, shape, device, name, dtype=torch.uint8):
    if out is None:
        return torch.empty(shape, dtype=dtype, device=device)
    _require_cuda(out, name, dtype)
    if tuple(out.shape) != tuple(shape):
        raise ValueError(f"{name}: expected shape {tuple(shape)}, got {tuple(out.shape)}")
    return out


INDEX_MAX_OPS = 128      # frames with longer draw lists are culled on the device instead


def build_overlay_index(ops_per_frame, width):
    """Spatial index of a draw list for frames ``width`` pixels wide (layout: include/aiko_b200.h).
    -> (uint32 ndarray [frames, frame_words], segments, words) or None when a list is too long."""
    most = max((len(o) for o in ops_per_frame), default=0)
    if most == 0 or most > INDEX_MAX_OPS:
        return None
    words = (most + 31) // 32
    segments = ((width + 3) // 4 + 31) // 32
    lane_lo = (np.arange(segments)[:, None] * 128 + np.arange(32)[None, :] * 4)          # [seg, lane]
    lane_hi = lane_lo + 3
    frame_words = 2 * segments * words * 32 + words * 32 * 2
    out = np.zeros((len(ops_per_frame), frame_words), np.uint32)
    cache = {}
    for f, frame_ops in enumerate(ops_per_frame):
        key = id(frame_ops)
        if key in cache:                     # the same list object for many frames (one overlay per batch)
            out[f] = out[cache[key]]
            continue
        cache[key] = f
        n = len(frame_ops)
        if n == 0:
            continue
        kind = np.array([o.kind for o in frame_ops])
        x0 = np.array([o.x0 for o in frame_ops]); x1 = np.array([o.x1 for o in frame_ops])
        y0 = np.array([o.y0 for o in frame_ops]); y1 = np.array([o.y1 for o in frame_ops])
        rect = kind == 0
        grow = rect.astype(np.int64)
        # side variant: two intervals A, B (equal for glyphs); band variant: the whole span
        a0 = x0 - grow; a1 = np.where(rect, x0 + 1, x1); b0 = np.where(rect, x1 - 1, x0); b1 = x1 + grow
        full0, full1 = x0 - grow, x1 + grow

        def touches(lo, hi):                 # [op, seg, lane]
            return (lo[:, None, None] <= lane_hi[None]) & (hi[:, None, None] >= lane_lo[None])

        side = touches(a0, a1) | touches(b0, b1)
        band = touches(full0, full1)
        xm = np.zeros((2, segments, words, 32), np.uint32)
        for b in range(n):
            w, bit = divmod(b, 32)
            xm[0, :, w, :] |= (side[b].astype(np.uint32) << np.uint32(bit))
            xm[1, :, w, :] |= (band[b].astype(np.uint32) << np.uint32(bit))
        clip = lambda v: np.clip(v, -32768, 32767).astype(np.int64) & 0xffff
        yinfo = np.zeros((words * 32, 2), np.uint32)
        yinfo[:, 0] = 1                      # unused slots: first_row 1 > last_row 0 -> never hit
        yinfo[:n, 0] = (clip(y0 - grow) | (clip(y1 + grow) << 16)).astype(np.uint32)
        yb0 = np.where(rect, y0 + 1, 32767); yb1 = np.where(rect, y1 - 1, -32768)
        yinfo[:n, 1] = (clip(yb0) | (clip(yb1) << 16)).astype(np.uint32)
        out[f] = np.concatenate([xm.reshape(-1), yinfo.reshape(-1)])
    return out, segments, words


GROUPS_PER_CTA = 256


def build_group_list(ops_per_frame, width, height):
    """The 4-pixel groups (pixels 4g .. 4g+3 of a row) that each frame's draw list can touch, with
    a bit mask of the ops per group, arranged for one thread per (frame, touched group)
    (layout: include/aiko_b200.h).  Frames that share one list object share their records.
    -> (records uint32 [n_records, 1 + words], ctas uint32 [n_ctas, 4], words) or None."""
    most = max((len(o) for o in ops_per_frame), default=0)
    if most == 0 or most > INDEX_MAX_OPS or sum(len(o) for o in ops_per_frame) >= (1 << 24):
        return None
    words = (most + 31) // 32
    groups = (width + 3) // 4
    first_ops = np.concatenate([[0], np.cumsum([len(o) for o in ops_per_frame])])
    cache, records, ctas, n_records = {}, [], [], 0
    for f, frame_ops in enumerate(ops_per_frame):
        key = id(frame_ops)
        if key not in cache:
            mask = np.zeros((height, groups, words), np.uint32)

            def mark(ya, yb, xa, xb, w, bit):
                ya, yb, xa, xb = max(ya, 0), min(yb, height - 1), max(xa, 0), min(xb, width - 1)
                if ya <= yb and xa <= xb:
                    mask[ya:yb + 1, xa // 4:xb // 4 + 1, w] |= np.uint32(1 << bit)

            for b_, o in enumerate(frame_ops):
                w, bit = divmod(b_, 32)
                if o.kind == 0:        # rectangle outline: top / bottom bands, then the two side bands
                    mark(o.y0 - 1, o.y0 + 1, o.x0 - 1, o.x1 + 1, w, bit)
                    mark(o.y1 - 1, o.y1 + 1, o.x0 - 1, o.x1 + 1, w, bit)
                    mark(o.y0 + 2, o.y1 - 2, o.x0 - 1, o.x0 + 1, w, bit)
                    mark(o.y0 + 2, o.y1 - 2, o.x1 - 1, o.x1 + 1, w, bit)
                else:
                    mark(o.y0, o.y1, o.x0, o.x1, w, bit)
            ys, gs = np.nonzero(mask.any(axis=2))
            rec = np.empty((len(ys), 1 + words), np.uint32)
            rec[:, 0] = gs.astype(np.uint32) | (ys.astype(np.uint32) << 16)
            rec[:, 1:] = mask[ys, gs]
            cache[key] = (n_records, len(ys))
            records.append(rec)
            n_records += len(ys)
        offset, count = cache[key]
        for start in range(0, count, GROUPS_PER_CTA):
            ctas.append((offset + start, min(GROUPS_PER_CTA, count - start), f, first_ops[f] | (len(frame_ops) << 24)))
    if not ctas:
        return None
    return np.concatenate(records), np.array(ctas, np.uint32), words


class DeviceOverlay:
    """A draw list resident on the GPU (see aiko_b200_overlay in include/aiko_b200.h)."""

    def __init__(self, ops, op_offsets, atlas=None, max_ops_per_frame=0, index=None, index_width=0, groups=None,
                 index_height=0):
        self.ops, self.op_offsets, self.atlas = ops, op_offsets, atlas
        self.struct = Overlay()
        self.struct.max_ops_per_frame = int(max_ops_per_frame)
        self.index = None
        if index is not None:
            table, segments, words = index
            self.index = table
            self.struct.index = table.data_ptr()
            self.struct.index_width, self.struct.index_segments = int(index_width), int(segments)
            self.struct.index_words, self.struct.index_frame_words = int(words), int(table.shape[1])
        self.groups = None
        if groups is not None and index_width:
            records, ctas, words = groups
            self.groups = (records, ctas)
            self.struct.index_width, self.struct.index_height = int(index_width), int(index_height)
            self.struct.group_records, self.struct.group_ctas = records.data_ptr(), ctas.data_ptr()
            self.struct.n_group_ctas, self.struct.group_words = int(ctas.shape[0]), int(words)
            if ctas.shape[0] == 0:             # "nothing is drawn": the pointers only need to be non-null
                self.struct.group_ctas = records.data_ptr()
        self.struct.ops = ops.data_ptr()
        self.struct.op_offsets = op_offsets.data_ptr()
        if atlas is not None:
            self.struct.patch_ids = atlas.patch_ids.data_ptr()
            self.struct.luts = atlas.luts.data_ptr()
            self.struct.patch_w, self.struct.patch_h = atlas.patch_w, atlas.patch_h
            self.struct.n_patches = atlas.patch_ids.shape[0]

    @staticmethod
    def pack_host(ops_per_frame):
        """-> (uint8 ndarray of packed ops, int32 ndarray of offsets)"""
        flat = [op for ops in ops_per_frame for op in ops]
        offsets = np.zeros(len(ops_per_frame) + 1, np.int32)
        if ops_per_frame:
            np.cumsum([len(o) for o in ops_per_frame], out=offsets[1:])
        count = max(len(flat), 1)
        arr = (OverlayOp * count)(*flat)
        raw = np.frombuffer(arr, dtype=np.uint8, count=ctypes.sizeof(OverlayOp) * count).copy()
        return raw, offsets

    @staticmethod
    def from_host(ops_per_frame, device, atlas=None, width=None, height=None):
        """ops_per_frame: list (one entry per frame) of lists of OverlayOp; ``width`` / ``height``:
        frame size the ops were built for (enables the spatial indices the kernels use)."""
        raw, offsets = DeviceOverlay.pack_host(ops_per_frame)
        most = max((len(o) for o in ops_per_frame), default=0)
        index = None
        if width and os.environ.get("AIKO_B200_OVERLAY_FUSED"):      # only the single-kernel variant uses it
            built = build_overlay_index(ops_per_frame, width)
            if built is not None:
                index = (torch.from_numpy(built[0]).to(device), built[1], built[2])
        groups = None
        if width and height:
            if most == 0:                       # nothing to draw at all: no kernel is launched
                empty = torch.zeros((1, 4), dtype=torch.int32, device=device)
                groups = (empty, empty[:0], 1)
            else:
                built = build_group_list(ops_per_frame, width, height)
                if built is not None:
                    groups = (torch.from_numpy(built[0].view(np.int32)).to(device),
                              torch.from_numpy(built[1].view(np.int32)).to(device), built[2])
        return DeviceOverlay(torch.from_numpy(raw).to(device), torch.from_numpy(offsets).to(device), atlas, most,
                             index, width or 0, groups, height or 0)

    def pointer(self):
        return ctypes.pointer(self.struct)


def reads_source_rows_once(frame_shape, width, height):
    """True when ``resize_linear`` to (width, height) fetches every source row it needs exactly once
    with bulk row copies (RGB, 16-byte aligned rows, horizontal scale <= 4, vertical scale >= 2):
    the case in which the kernel may as well read the rows from page-locked host memory."""
    if len(frame_shape) != 3 or frame_shape[2] != 3:
        return False
    h, w = frame_shape[:2]
    return (w * 3) % 16 == 0 and w <= 4 * width and h >= 2 * height and width > 0 and height > 0


def resize_linear(src, width, height, *, swap_rb=False, overlay=None, out=None, stream=None, device=None):
    """cv2.resize(image, (width, height)) for every frame of ``src``; optional fused overlay / swap.

    ``src`` is a CUDA tensor, or -- with ``device`` given -- a page-locked host tensor that the
    kernel reads over PCIe itself (no staging copy in HBM; see ``reads_source_rows_once``).
    """
    if device is not None and isinstance(src, torch.Tensor) and not src.is_cuda:
        if not src.is_pinned() or src.dtype != torch.uint8 or not src.is_contiguous():
            raise RuntimeError("resize_linear src: a host tensor must be page-locked, uint8 and contiguous")
        dev = torch.device(device)
    else:
        _require_cuda(src, "resize_linear src")
        dev = src.device
    n, h, w, c, gray = _nhwc(src)
    shape = (n, height, width) if gray else (n, height, width, c)
    out = _out(out, shape, dev, "resize_linear out")
    with torch.cuda.device(dev):
        check(_native.lib().aiko_b200_resize_linear_u8(
            src.data_ptr(), n, h, w, c, w * c, h * w * c, out.data_ptr(), height, width, width * c,
            height * width * c, int(bool(swap_rb)), overlay.pointer() if overlay is not None else None,
            _stream_ptr(stream)))
    return out


_bicubic_plans = {}   # (device, sh, sw, dh, dw) -> prepared workspace tensor


def bicubic_workspace(device, sh, sw, dh, dw, stream=None):
    """Prepared weight tables for PIL-style resizing of sh x sw frames to dh x dw (cached per device)."""
    key = (str(device), sh, sw, dh, dw)
    ws = _bicubic_plans.get(key)
    if ws is None:
        nbytes = _native.lib().aiko_b200_resize_bicubic_workspace_bytes(sh, sw, dh, dw)
        if nbytes == 0:
            raise ValueError(f"resize_bicubic_aa: bad size {sw}x{sh} -> {dw}x{dh}")
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        with torch.cuda.device(device):
            check(_native.lib().aiko_b200_resize_bicubic_prepare(ws.data_ptr(), nbytes, sh, sw, dh, dw,
                                                                 _stream_ptr(stream)))
        # later calls may come from other streams: make the tables visible to all of them
        (stream or torch.cuda.current_stream()).synchronize()
        _bicubic_plans[key] = ws
    return ws


def resize_bicubic_aa(src, width, height, *, out=None, stream=None):
    """PIL.Image.resize((width, height)) (default bicubic, antialiased) for every frame of ``src``."""
    _require_cuda(src, "resize_bicubic_aa src")
    n, h, w, c, gray = _nhwc(src)
    shape = (n, height, width) if gray else (n, height, width, c)
    out = _out(out, shape, src.device, "resize_bicubic_aa out")
    if n == 0:
        return out
    ws = bicubic_workspace(src.device, h, w, height, width, stream)
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_resize_bicubic_aa_u8(src.data_ptr(), n, h, w, c, out.data_ptr(), height, width,
                                                           ws.data_ptr(), ws.numel(), _stream_ptr(stream)))
    return out


def resize_norm_f32(src, width, height, mean, std, *, out=None, stream=None):
    """cv2-exact bilinear resize of RGB frames, then ((x / 255) - mean) / std as float32 [N, 3, H, W]."""
    _require_cuda(src, "resize_norm_f32 src")
    n, h, w, c, _ = _nhwc(src)
    if c != 3:
        raise ValueError("resize_norm_f32: RGB frames expected")
    out = _out(out, (n, 3, height, width), src.device, "resize_norm_f32 out", torch.float32)
    m = (ctypes.c_float * 3)(*[float(v) for v in mean])
    s = (ctypes.c_float * 3)(*[float(v) for v in std])
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_resize_norm_f32_nchw(src.data_ptr(), n, h, w, out.data_ptr(), height, width,
                                                           m, s, _stream_ptr(stream)))
    return out


def overlay_draw(src, overlay, *, out=None, stream=None):
    """ImageOverlay's drawing on frames already resident on the GPU."""
    _require_cuda(src, "overlay src")
    n, h, w, c, _ = _nhwc(src)
    out = _out(out, src.shape, src.device, "overlay out")
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_overlay_u8(src.data_ptr(), out.data_ptr(), n, h, w, c, w * c, h * w * c,
                                                 overlay.pointer(), _stream_ptr(stream)))
    return out


def swap_rb(src, *, out=None, stream=None):
    _require_cuda(src, "swap_rb src")
    if src.shape[-1] != 3:
        raise ValueError("swap_rb: last dimension must be 3")
    out = _out(out, src.shape, src.device, "swap_rb out")
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_swap_rb_u8(src.data_ptr(), out.data_ptr(), src.numel() // 3,
                                                 _stream_ptr(stream)))
    return out


def bgr2gray(src, *, rgb_order=False, out=None, stream=None):
    _require_cuda(src, "bgr2gray src")
    if src.shape[-1] != 3:
        raise ValueError("bgr2gray: last dimension must be 3")
    out = _out(out, src.shape[:-1], src.device, "bgr2gray out")
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_bgr2gray_u8(src.data_ptr(), out.data_ptr(), src.numel() // 3,
                                                  int(bool(rgb_order)), _stream_ptr(stream)))
    return out


def gray2bgr(src, *, out=None, stream=None):
    _require_cuda(src, "gray2bgr src")
    out = _out(out, tuple(src.shape) + (3,), src.device, "gray2bgr out")
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_gray2bgr_u8(src.data_ptr(), out.data_ptr(), src.numel(),
                                                  _stream_ptr(stream)))
    return out


def flip(src, flip_code, *, out=None, stream=None):
    """cv2.flip semantics: 0 vertical, >0 horizontal, <0 both."""
    _require_cuda(src, "flip src")
    n, h, w, c, _ = _nhwc(src)
    out = _out(out, src.shape, src.device, "flip out")
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_flip_u8(src.data_ptr(), out.data_ptr(), n, h, w, c, int(flip_code),
                                              _stream_ptr(stream)))
    return out


def crop(src, x0, y0, width, height, *, out=None, stream=None):
    _require_cuda(src, "crop src")
    n, h, w, c, gray = _nhwc(src)
    shape = (n, height, width) if gray else (n, height, width, c)
    out = _out(out, shape, src.device, "crop out")
    with torch.cuda.device(src.device):
        check(_native.lib().aiko_b200_crop_u8(src.data_ptr(), n, h, w, c, out.data_ptr(), int(x0), int(y0),
                                              int(width), int(height), _stream_ptr(stream)))
    return out
