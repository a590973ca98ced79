"""Python-facing operators over the C ABI.

torch is used here for device memory and streams only (``tensor.data_ptr()``,
``torch.cuda.current_stream()``); the arithmetic is in libaiko_b200.so.  Each function
mirrors the library call the reference element makes (cited in include/aiko_b200.h) but
takes a whole batch ``uint8[N, H, W, C]`` (or ``[N, H, W]`` for grayscale) resident on the GPU.
"""
import contextlib
import ctypes
from array import array

import numpy as np
import torch

from . import _native
from ._native import Overlay, OverlayOp, check


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


def _stream_int(stream=None):
    return (stream if stream is not None else torch.cuda.current_stream()).cuda_stream


def _T():
    """torch.ops.aiko_b200: the C ABI as PyTorch custom ops (torch_ops.py)."""
    global _T_cache
    if _T_cache is None:
        from . import torch_ops
        _T_cache = torch_ops.OPS
    return _T_cache


_T_cache = None


class _on_device:
    """``torch.cuda.device(dev)`` that costs nothing when ``dev`` already is the current device."""
    __slots__ = ("ctx",)

    def __init__(self, dev):
        index = dev.index if dev.index is not None else torch.cuda.current_device()
        self.ctx = None if index == torch.cuda.current_device() else torch.cuda.device(index)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            return self.ctx.__exit__(*a)
        return False


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


class DeviceOverlay:
    """Draw lists resident on the GPU (see aiko_b200_overlay in include/aiko_b200.h).  One device
    buffer holds ops | op_offsets | list_of_frame, uploaded with a single copy."""

    def __init__(self, buffer, n_ops, n_lists, n_frames, ops_bytes, atlas, width, height, max_ops, max_marks,
                 max_cells):
        self.buffer, self.atlas, self.n_ops = buffer, atlas, n_ops
        st = self.struct = Overlay()
        base = buffer.data_ptr()
        st.ops, st.op_offsets = base, base + ops_bytes
        st.list_of_frame = base + ops_bytes + 4 * (n_lists + 1)
        st.n_lists, st.n_frames = int(n_lists), int(n_frames)
        st.max_ops_per_list, st.max_marks_per_list = int(max_ops), int(max_marks)
        st.max_cells_per_list = int(max_cells)
        st.frame_width, st.frame_height = int(width or 0), int(height or 0)
        if atlas is not None:
            st.patch_ids, st.luts = atlas.patch_ids.data_ptr(), atlas.luts.data_ptr()
            st.patch_w, st.patch_h = atlas.patch_w, atlas.patch_h
            st.n_patches = atlas.patch_ids.shape[0]

    @staticmethod
    def from_plan(plan, device, atlas=None, pinned=None, stream=None):
        """``plan``: overlay_plan.Plan.  ``pinned``: optional pinned uint8 tensor the plan was written
        into (then the copy is asynchronous; the caller keeps it untouched until the copy is done)."""
        device = torch.device(device)
        if atlas is None:
            from .text_atlas import TextAtlas
            atlas = TextAtlas.shared().device(device)
        s = stream if stream is not None else torch.cuda.current_stream(device)
        if atlas.ready is not None:
            s.wait_event(atlas.ready)            # the tables may have been appended to on another stream
        n = plan.buffer.nbytes
        with torch.cuda.stream(s):
            buf = torch.empty(n, dtype=torch.uint8, device=device)
            if pinned is not None and plan.buffer.ctypes.data == pinned.data_ptr():
                buf.copy_(pinned[:n], non_blocking=True)
            else:
                buf.copy_(torch.from_numpy(plan.buffer))
        i = plan.info
        return DeviceOverlay(buf, plan.n_ops, plan.n_lists, plan.n_frames, plan.ops_bytes, atlas, plan.width,
                             plan.height, i.max_ops_per_list, i.max_marks_per_list, i.max_cells_per_list)

    @staticmethod
    def from_host(ops_per_frame, device, atlas=None, width=None, height=None):
        """ops_per_frame: list (one entry per frame) of lists of OverlayOp made by other means than
        ``overlay_plan.plan``; frames that pass the same list object share it.  ``width`` / ``height``:
        the frame size (lets the library run one thread per touched 4-pixel group)."""
        lists, list_of_frame, seen = [], [], {}
        for o in ops_per_frame:
            li = seen.get(id(o))
            if li is None:
                li = seen[id(o)] = len(lists)
                lists.append(o)
            list_of_frame.append(li)
        flat = [op for o in lists for op in o]
        offsets = np.zeros(len(lists) + 1, np.int32)
        if lists:
            np.cumsum([len(o) for o in lists], out=offsets[1:])
        count = max(len(flat), 1)
        arr = (OverlayOp * count)(*flat)
        max_ops = max((len(o) for o in lists), default=0)
        max_marks = max_cells = 0
        if width and height:
            marks, cells = ctypes.c_int(), ctypes.c_longlong()
            for i in range(len(lists)):
                first = int(offsets[i])
                check(_native.lib().aiko_b200_overlay_cells_host(
                    ctypes.byref(arr, first * ctypes.sizeof(OverlayOp)), len(lists[i]), int(width), int(height),
                    ctypes.byref(marks), ctypes.byref(cells)))
                max_marks, max_cells = max(max_marks, marks.value), max(max_cells, min(cells.value, 2 ** 31 - 1))
        ops_bytes = count * 32
        raw = np.concatenate([np.frombuffer(arr, dtype=np.uint8, count=ops_bytes), offsets.view(np.uint8),
                              np.asarray(list_of_frame, np.int32).view(np.uint8)])
        device = torch.device(device)
        if atlas is not None and atlas.ready is not None:
            torch.cuda.current_stream(device).wait_event(atlas.ready)
        return DeviceOverlay(torch.from_numpy(raw).to(device), len(flat), len(lists), len(list_of_frame), ops_bytes,
                             atlas, width, height, max_ops, max_marks, max_cells)

    def pointer(self):
        return ctypes.pointer(self.struct)


class OverlayRing:
    """Draw lists that change with every call (detections move, label text carries the confidence):
    a ring of page-locked + device buffer pairs allocated once.  A new list costs one host planning
    call; its bytes travel inside the call that draws it (``upload_from_host`` of aiko_b200_overlay),
    so there is no allocation, no extra copy call and no synchronisation unless the ring wraps onto
    copies still in flight.  The DeviceOverlay returned by ``plan`` must be used (once, on
    ``stream``) before the next ``plan`` and stays valid for ``slots`` / 2 further calls."""

    def __init__(self, device, capacity=1 << 16, slots=16):
        from .text_atlas import TextAtlas
        self.device = torch.device(device)
        self.atlas = TextAtlas.shared()
        self.capacity = capacity
        self.pinned = torch.empty((slots, capacity), dtype=torch.uint8).pin_memory()
        self.dev = torch.empty((slots, capacity), dtype=torch.uint8, device=self.device)
        self.views = [self.pinned[i].numpy() for i in range(slots)]
        self.overlays = [None] * slots
        self.half = max(slots // 2, 1)
        self.events = [None, None]        # one per half of the ring: recorded after the half's last use
        self.next = 0
        self._atlas_dev = None
        self._info = _native.OverlayPlanInfo()

    def _slot(self, s):
        """Next ring slot (waits only when the ring wraps onto work still in flight) -> its DeviceOverlay."""
        i = self.next
        n = len(self.views)
        if i == 0 or i == self.half:
            # entering a half: everything queued from it a lap ago must have run (the event recorded when
            # the other half was entered); and everything queued so far covers the other half's last lap
            h = 0 if i == 0 else 1
            if n == 1:
                s.synchronize()
            else:
                if self.events[h] is not None:
                    self.events[h].synchronize()
                if self.events[1 - h] is None:
                    self.events[1 - h] = torch.cuda.Event()
                self.events[1 - h].record(s)
        self.next = (i + 1) % n
        ov = self.overlays[i]
        if ov is None:
            atlas = self._atlas_dev or self.atlas.device(self.device)
            ov = self.overlays[i] = DeviceOverlay(self.dev[i], 0, 0, 0, 0, atlas, 0, 0, 0, 0, 0)
            ov.host_ptr, ov.dev_ptr = self.pinned[i].data_ptr(), self.dev[i].data_ptr()
        return ov

    def _sync_atlas(self, ov, info, s):
        """Bring the device copy of the glyph atlas to the state the planner left it in."""
        atlas = self._atlas_dev
        if atlas is None or atlas.version != info.atlas_version or atlas.epoch != info.atlas_epoch:
            atlas = self._atlas_dev = self.atlas.device(self.device)       # uploads what the planner appended
            if atlas.ready is not None:
                s.wait_event(atlas.ready)
        if ov.atlas is not atlas:                        # a new atlas epoch: the tables moved
            ov.atlas = atlas
            ov.struct.patch_ids, ov.struct.luts = atlas.patch_ids.data_ptr(), atlas.luts.data_ptr()
        ov.n_ops = info.n_ops

    def plan(self, overlays, width, height, layout="rgb", n_frames=None, stream=None):
        from .overlay_plan import LAYOUTS, gather_objects, plan
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        ov = self._slot(s)
        n_lists, offsets, rects, labels, list_of_frame = gather_objects(overlays)
        if list_of_frame is None:
            frames, frame_map, keep = int(n_frames or 1), None, None
        else:
            keep = array("i", list_of_frame)               # (stays alive until the call has returned)
            frames, frame_map = len(list_of_frame), keep.buffer_info()[0]
        info = self._info
        rc = _native.lib().aiko_b200_overlay_plan_struct_host(
            n_lists, offsets.buffer_info()[0], rects.buffer_info()[0], labels, int(width), int(height),
            LAYOUTS[layout], frames, frame_map, ov.host_ptr, self.capacity, ov.dev_ptr, ov.struct, info)
        if rc == _native.ERR_WORKSPACE:                  # does not fit a slot: one-off buffers
            planned = plan(overlays, width, height, layout, n_frames=n_frames)
            return DeviceOverlay.from_plan(planned, self.device, self.atlas.device(self.device), stream=s)
        check(rc)
        self._sync_atlas(ov, info, s)
        return ov

    def resize_draw(self, src, overlays, width, height, *, swap_rb=False, out=None, stream=None):
        """``resize_linear(src, width, height, swap_rb=..., overlay=self.plan(overlays, ...))`` as one call of
        the library (aiko_b200_resize_linear_objects_u8): ImageResize -> ImageOverlay (-> RGB2BGR) for
        detections that are new on every call.  ``overlays``: one dict for all frames or one per frame;
        colours follow the stored channel order (BGR when ``swap_rb``)."""
        from .overlay_plan import gather_objects
        _require_cuda(src, "resize_draw src")
        n, h, w, c, gray = _nhwc(src)
        out = _out(out, (n, height, width) if gray else (n, height, width, c), src.device, "resize_draw out")
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        ov = self._slot(s)
        n_lists, offsets, rects, labels, list_of_frame = gather_objects(overlays)
        if list_of_frame is not None and len(list_of_frame) != n:
            raise ValueError(f"resize_draw: {len(list_of_frame)} overlays for {n} frames")
        keep = array("i", list_of_frame) if list_of_frame is not None else None     # alive across the call
        frame_map = keep.buffer_info()[0] if keep is not None else None
        info, atlas = self._info, self._atlas_dev
        if atlas is None:
            atlas = self._atlas_dev = self.atlas.device(self.device)
            if atlas.ready is not None:
                s.wait_event(atlas.ready)
        if ov.atlas is not atlas:
            ov.atlas = atlas
            ov.struct.patch_ids, ov.struct.luts = atlas.patch_ids.data_ptr(), atlas.luts.data_ptr()
        lib = _native.lib()
        with _on_device(self.device):
            rc = lib.aiko_b200_resize_linear_objects_u8(
                src.data_ptr(), n, h, w, c, w * c, h * w * c, out.data_ptr(), height, width, width * c,
                height * width * c, int(swap_rb), n_lists, offsets.buffer_info()[0], rects.buffer_info()[0], labels,
                1 if swap_rb else 0, frame_map, ov.host_ptr, self.capacity, ov.dev_ptr, ov.struct, info,
                atlas.version, atlas.epoch, s.cuda_stream)
            if rc == _native.ERR_WORKSPACE or rc == _native.ATLAS_GREW:
                # the lists do not fit a ring slot, or the planner appended glyph variants: the two-call path
                if rc == _native.ATLAS_GREW:
                    self._sync_atlas(ov, info, s)
                else:
                    from .overlay_plan import plan
                    planned = plan(overlays, width, height, "bgr" if swap_rb else "rgb", n_frames=n)
                    ov = DeviceOverlay.from_plan(planned, self.device, self.atlas.device(self.device), stream=s)
                rc = lib.aiko_b200_resize_linear_u8(
                    src.data_ptr(), n, h, w, c, w * c, h * w * c, out.data_ptr(), height, width, width * c,
                    height * width * c, int(swap_rb), ov.struct if ov.n_ops else None, s.cuda_stream)
        check(rc)
        return out


class DeviceDetections:
    """A detector's output as it sits on the GPU, in place of the overlay dict the reference builds on the host
    (R/examples/yolo/yolo.py:56-87): ``boxes`` int32 [L, K, 4] (x, y, w, h), ``class_ids`` int32 [L, K],
    ``confidences`` float32 [L, K]; L = 1 (one overlay for all frames of the call, as in the reference) or one
    list per frame; unused slots carry confidence 0.  ``names``: class names (index = class id)."""

    _name_tables = {}

    def __init__(self, boxes, class_ids, confidences, names):
        for t, dtype, what in ((boxes, torch.int32, "boxes"), (class_ids, torch.int32, "class_ids"),
                               (confidences, torch.float32, "confidences")):
            if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == dtype and t.is_contiguous()):
                raise RuntimeError(f"DeviceDetections {what}: a contiguous {dtype} CUDA tensor is needed")
        if boxes.dim() == 2:
            boxes, class_ids, confidences = boxes[None], class_ids[None], confidences[None]
        if boxes.dim() != 3 or boxes.shape[2] != 4 or class_ids.shape != boxes.shape[:2] or confidences.shape != boxes.shape[:2]:
            raise ValueError("DeviceDetections: boxes [L, K, 4], class_ids [L, K], confidences [L, K]")
        self.boxes, self.class_ids, self.confidences, self.names = boxes, class_ids, confidences, tuple(names)

    def name_table(self, device):
        """(device char table [n_names, stride], stride), cached per (names, device)."""
        from .text_atlas import _encode
        key = (self.names, str(device))
        entry = self._name_tables.get(key)
        if entry is None:
            raw = [_encode(n) if n else b"" for n in self.names]
            stride = max([len(r) for r in raw] + [1]) + 1
            if stride > 40:
                raise ValueError("DeviceDetections: class names longer than 39 characters")
            table = torch.zeros((max(len(raw), 1), stride), dtype=torch.uint8)
            for k, r in enumerate(raw):
                table[k, :len(r)] = torch.frombuffer(bytearray(r), dtype=torch.uint8) if r else table[k, :0]
            entry = self._name_tables[key] = (table.to(device), stride)
        return entry

    def to_overlay_dicts(self):
        """The overlay dict(s) the reference would have built (reads the tensors back: the host path)."""
        boxes, cls, conf = self.boxes.cpu().numpy(), self.class_ids.cpu().numpy(), self.confidences.cpu().numpy()
        dicts = []
        for li in range(boxes.shape[0]):
            objects, rectangles = [], []
            for k in range(boxes.shape[1]):
                c = int(cls[li, k])
                name = self.names[c] if 0 <= c < len(self.names) and self.names[c] else None
                objects.append({"name": name, "confidence": float(conf[li, k])})
                x, y, w, h = (int(v) for v in boxes[li, k])
                rectangles.append({"x": x, "y": y, "w": w, "h": h})
            dicts.append({"objects": objects, "rectangles": rectangles})
        return dicts


_boxes_workspaces = {}


def overlay_boxes(frames, detections, *, layout="rgb", out=None, stream=None, check=True):
    """ImageOverlay drawn from a detector's device-resident output (``DeviceDetections``) on RGB frames
    [N, H, W, 3]: planning and drawing both run on the GPU (aiko_b200_overlay_boxes_u8).  In place unless ``out``
    is given.  A label that comes within the border margin of the frame needs a glyph variant rasterised against
    that border; the variants the host has made so far are mirrored on the device.  When one is missing, nothing is
    drawn and the status says so: with ``check`` (the default) the 4-byte status is read back and the call is redone
    through the host planner (which rasterises the variant: the next call finds it on the device); without it the
    status tensor is returned next to the frames for the caller to look at later.  -> frames, or (frames, status) when ``check`` is false."""
    from .overlay_plan import LAYOUTS
    from .text_atlas import TextAtlas
    _require_cuda(frames, "overlay_boxes frames")
    if frames.dim() != 4 or frames.shape[3] != 3:
        raise ValueError("overlay_boxes: RGB frames [N, H, W, 3] are needed")
    n, h, w, _ = frames.shape
    d = detections
    lists, slots = d.boxes.shape[0], d.boxes.shape[1]
    if lists not in (1, n):
        raise ValueError(f"overlay_boxes: {lists} lists for {n} frames (1, or one per frame)")
    dev = frames.device
    s = stream if stream is not None else torch.cuda.current_stream(dev)
    if out is not None:
        out = _out(out, frames.shape, dev, "overlay_boxes out")
        if out.data_ptr() != frames.data_ptr():
            with torch.cuda.stream(s):
                out.copy_(frames)
        frames = out
    lib = _native.lib()
    key = (str(dev), lists, slots, n)
    entry = _boxes_workspaces.get(key)
    if entry is None:
        nbytes = lib.aiko_b200_overlay_boxes_workspace_bytes(lists, slots, n)
        entry = _boxes_workspaces[key] = (torch.empty(nbytes, dtype=torch.uint8, device=dev),
                                          torch.zeros(1, dtype=torch.int32, device=dev))
    workspace, status = entry
    names, stride = d.name_table(dev)
    shared = TextAtlas.shared()
    atlas = shared.device(dev)
    if atlas.ready is not None:
        s.wait_event(atlas.ready)
    with torch.cuda.stream(s):
        keys, firsts, variant_boxes, n_variants = shared.variants(dev)     # (a copy only when the host made new variants)
    _T().overlay_boxes_u8(frames, LAYOUTS[layout], d.boxes, d.class_ids, d.confidences, names, len(d.names), atlas.patch_ids,
                          atlas.luts, atlas.n_patches, keys, firsts, variant_boxes, n_variants, workspace, status,
                          s.cuda_stream)
    if not check:
        return frames, status
    with torch.cuda.stream(s):
        code = int(status.cpu()[0])
    if code != 0:
        # border-clipped labels (or an unusual value): this call through the host planner, boxes read back
        from .overlay_plan import plan
        dicts = d.to_overlay_dicts()
        planned = plan(dicts[0] if lists == 1 else dicts, w, h, layout, n_frames=n)
        ov = DeviceOverlay.from_plan(planned, dev, TextAtlas.shared().device(dev), stream=s)
        overlay_draw(frames, ov, out=frames, stream=s)
    return frames


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
    if overlay is not None and overlay.n_ops == 0:      # nothing to draw
        overlay = None
    _T().resize_linear_u8(src, out, bool(swap_rb), ctypes.addressof(overlay.struct) if overlay is not None else 0,
                          _stream_int(stream))
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
            _T().resize_bicubic_prepare(ws, sh, sw, dh, dw, _stream_int(stream))
        # later calls may come from other streams: make the tables visible to all of them
        (stream or torch.cuda.current_stream()).synchronize()
        if nbytes >= 8 and int(ws[:8].view(torch.int32)[1]) != 0:      # set by prepare when a table overflowed
            raise _native.NativeError(_native.ERR_UNSUPPORTED,
                                      f"resize_bicubic_aa: weight tables cannot hold {sw}x{sh} -> {dw}x{dh}")
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
    with _on_device(src.device):
        _T().resize_bicubic_aa_u8(src, out, ws, _stream_int(stream))
    return out


PIL_PLAIN_MODES = ("L", "RGB", "CMYK", "YCbCr", "HSV", "LAB")     # 8-bit bands, filtered band by band
PIL_ALPHA_MODES = ("LA", "RGBA")                                  # premultiplied around the filter
PIL_NEAREST_MODES = ("P", "1")                                    # Image.resize forces NEAREST
_nearest_tables = {}


def premultiply(src, *, inverse=False, channels=None, out=None, stream=None):
    """Pillow's RGBA -> RGBa / LA -> La conversion (and back with ``inverse``).  Forward: ``src``
    [..., 2 or 4] -> [..., 4]; inverse: ``src`` [..., 4] -> [..., channels]."""
    _require_cuda(src, "premultiply src")
    c = int(channels if inverse else src.shape[-1])
    shape = tuple(src.shape[:-1]) + ((c,) if inverse else (4,))
    out = _out(out, shape, src.device, "premultiply out")
    with _on_device(src.device):
        _T().premultiply_u8(src, out, c, bool(inverse), _stream_int(stream))
    return out


def _nearest_table(src, dst):
    """Source index per destination index as Pillow's ImagingScaleAffine tabulates it: the position
    starts at 0.5 * (src / dst) and grows by src / dst per step, summed sequentially in double."""
    a = src / dst
    pos = np.add.accumulate(np.concatenate([[a * 0.5], np.full(dst - 1, a)]))
    idx = np.where(pos < 0.0, -1, pos.astype(np.int64))
    return np.where((idx >= 0) & (idx < src), idx, -1).astype(np.int32)


def resize_nearest_pil(src, width, height, *, out=None, stream=None):
    """PIL.Image.resize((width, height), NEAREST) -- what Pillow does for modes P and 1."""
    _require_cuda(src, "resize_nearest_pil src")
    n, h, w, c, gray = _nhwc(src)
    shape = (n, height, width) if gray else (n, height, width, c)
    out = _out(out, shape, src.device, "resize_nearest_pil out")
    key = (str(src.device), h, w, height, width)
    tabs = _nearest_tables.get(key)
    if tabs is None:
        tabs = _nearest_tables[key] = (torch.from_numpy(_nearest_table(w, width)).to(src.device),
                                       torch.from_numpy(_nearest_table(h, height)).to(src.device))
        torch.cuda.current_stream(src.device).synchronize()
    with _on_device(src.device):
        _T().gather_nearest_u8(src, out, tabs[0], tabs[1], _stream_int(stream))
    return out


def resize_pil(src, mode, width, height, *, stream=None):
    """PIL.Image.resize((width, height)) for frames of PIL mode ``mode`` (uint8 [N, H, W(, bands)]):
    antialiased bicubic per band; RGBA / LA premultiplied around it; P / 1 nearest."""
    if mode in PIL_PLAIN_MODES:
        return resize_bicubic_aa(src, width, height, stream=stream)
    if mode in PIL_ALPHA_MODES:
        pre = premultiply(src, stream=stream)
        return premultiply(resize_bicubic_aa(pre, width, height, stream=stream), inverse=True,
                           channels=src.shape[-1], stream=stream)
    if mode in PIL_NEAREST_MODES:
        return resize_nearest_pil(src, width, height, stream=stream)
    raise ValueError(f"PIL mode {mode!r} is not supported (8-bit modes only)")


def resize_norm_f32(src, width, height, mean, std, *, out=None, stream=None):
    """cv2-exact bilinear resize of RGB frames, then ((x / 255) - mean) / std as float32 [N, 3, H, W]."""
    _require_cuda(src, "resize_norm_f32 src")
    n, h, w, c, _ = _nhwc(src)
    if c != 3:
        raise ValueError("resize_norm_f32: RGB frames expected")
    out = _out(out, (n, 3, height, width), src.device, "resize_norm_f32 out", torch.float32)
    with _on_device(src.device):
        _T().resize_norm_f32_nchw(src, out, [float(v) for v in mean], [float(v) for v in std], _stream_int(stream))
    return out


def overlay_draw(src, overlay, *, out=None, stream=None):
    """ImageOverlay's drawing on frames already resident on the GPU."""
    _require_cuda(src, "overlay src")
    n, h, w, c, _ = _nhwc(src)
    out = _out(out, src.shape, src.device, "overlay out")
    if overlay.n_ops == 0:                               # nothing to draw (the gray round trip is the identity)
        if out.data_ptr() != src.data_ptr():
            with torch.cuda.stream(stream) if stream is not None else contextlib.nullcontext():
                out.copy_(src)
        return out
    with _on_device(src.device):
        _T().overlay_u8(src, out, ctypes.addressof(overlay.struct), _stream_int(stream))
    return out


def swap_rb(src, *, out=None, stream=None):
    _require_cuda(src, "swap_rb src")
    if src.shape[-1] != 3:
        raise ValueError("swap_rb: last dimension must be 3")
    out = _out(out, src.shape, src.device, "swap_rb out")
    with _on_device(src.device):
        _T().swap_rb_u8(src, out, _stream_int(stream))
    return out


def bgr2gray(src, *, rgb_order=False, out=None, stream=None):
    _require_cuda(src, "bgr2gray src")
    if src.shape[-1] != 3:
        raise ValueError("bgr2gray: last dimension must be 3")
    out = _out(out, src.shape[:-1], src.device, "bgr2gray out")
    with _on_device(src.device):
        _T().bgr2gray_u8(src, out, bool(rgb_order), _stream_int(stream))
    return out


def yuv(src, *, bgr_order=False, inverse=False, out=None, stream=None):
    """cv2.cvtColor(RGB2YUV / BGR2YUV) of uint8 frames, or YUV2RGB / YUV2BGR with ``inverse``."""
    _require_cuda(src, "yuv src")
    if src.shape[-1] != 3:
        raise ValueError("yuv: last dimension must be 3")
    out = _out(out, src.shape, src.device, "yuv out")
    with _on_device(src.device):
        _T().yuv_u8(src, out, bool(bgr_order), bool(inverse), _stream_int(stream))
    return out


def alpha_paste(frames, rgba, x=0, y=0, *, out=None, stream=None):
    """PIL's ``frame.paste(image, (x, y), image)`` for every RGB frame of ``frames`` [N, H, W, 3]: the RGBA image
    ``rgba`` [h, w, 4] (CUDA tensor) blended over the frames by its own alpha.  In place unless ``out`` is given."""
    _require_cuda(frames, "alpha_paste frames")
    _require_cuda(rgba, "alpha_paste rgba")
    if frames.dim() != 4 or frames.shape[3] != 3 or rgba.dim() != 3 or rgba.shape[2] != 4:
        raise ValueError("alpha_paste: frames [N, H, W, 3] and an RGBA image [h, w, 4] are needed")
    if out is not None:
        out = _out(out, frames.shape, frames.device, "alpha_paste out")
        if out.data_ptr() != frames.data_ptr():
            with torch.cuda.stream(stream) if stream is not None else contextlib.nullcontext():
                out.copy_(frames)
        frames = out
    with _on_device(frames.device):
        _T().alpha_paste_u8(frames, rgba, int(x), int(y), _stream_int(stream))
    return frames


def gray2bgr(src, *, out=None, stream=None):
    _require_cuda(src, "gray2bgr src")
    out = _out(out, tuple(src.shape) + (3,), src.device, "gray2bgr out")
    with _on_device(src.device):
        _T().gray2bgr_u8(src, out, _stream_int(stream))
    return out


def flip(src, flip_code, *, out=None, stream=None):
    """cv2.flip semantics: 0 vertical, >0 horizontal, <0 both."""
    _require_cuda(src, "flip src")
    n, h, w, c, _ = _nhwc(src)
    out = _out(out, src.shape, src.device, "flip out")
    with _on_device(src.device):
        _T().flip_u8(src, out, int(flip_code), _stream_int(stream))
    return out


def crop(src, x0, y0, width, height, *, out=None, stream=None):
    _require_cuda(src, "crop src")
    n, h, w, c, gray = _nhwc(src)
    shape = (n, height, width) if gray else (n, height, width, c)
    out = _out(out, shape, src.device, "crop out")
    with _on_device(src.device):
        _T().crop_u8(src, out, int(x0), int(y0), _stream_int(stream))
    return out
