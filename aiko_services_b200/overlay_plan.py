"""ImageOverlay's per-object logic -> GPU draw lists.

The reference (R/elements/media/image_io.py:119-155) zips objects with rectangles, skips objects
at or below the confidence threshold, formats the label ``f"{name}: {confidence:0.2f}"``, places
it, and draws the label before its rectangle.  The dictionary handling and the text formatting
stay here (they are Python semantics); size, placement, glyphs and rectangles are planned by
``aiko_b200_overlay_plan_host`` in libaiko_b200.so (csrc/overlay_plan.cpp), which fills a caller
buffer with 32-byte ``OverlayOp`` records (include/aiko_b200.h) that the kernels replay in order.
"""
import ctypes
from array import array

import numpy as np

from . import _native
from ._native import OverlayOp, OverlayPlanInfo, check
from .text_atlas import TextAtlas, _encode

OP_RECT, OP_GLYPH = 0, 1
THRESHOLD = 0.0             # image_io.py:105
LAYOUTS = {"rgb": 0, "bgr": 1}     # channel order of the stored frame the ops will be applied to


def _objects(overlay, rects, labels):
    """image_io.py:127-139 for one overlay dict: appends x, y, w, h and the label bytes per drawn
    object; -> number of objects appended."""
    if "rectangles" not in overlay:
        return 0
    rectangles = overlay["rectangles"]
    objects = overlay["objects"] if "objects" in overlay else [{}] * len(rectangles)
    count = 0
    for obj, rectangle in zip(objects, rectangles):
        name = obj.get("name", None)
        confidence = obj.get("confidence", 1.0)
        if not confidence > THRESHOLD:
            continue
        rects += (int(rectangle["x"]), int(rectangle["y"]), int(rectangle["w"]), int(rectangle["h"]))
        labels.append(f"{name}: {confidence:0.2f}" if name else "")
        count += 1
    return count


class Plan:
    """Draw lists of one call, packed for a single host-to-device copy:
    ``buffer`` = ops (32 bytes each) | op_offsets int32[n_lists + 1] | list_of_frame int32[n_frames]."""

    __slots__ = ("buffer", "n_ops", "n_lists", "n_frames", "info", "width", "height", "ops_bytes")

    def ops(self):
        """The ops as a list of OverlayOp per list (tests, CPU replay)."""
        arr = (OverlayOp * max(self.n_ops, 1)).from_buffer_copy(self.buffer[:max(self.n_ops, 1) * 32].tobytes())
        offsets = self.buffer[self.ops_bytes:self.ops_bytes + 4 * (self.n_lists + 1)].view(np.int32)
        return [[arr[k] for k in range(offsets[i], offsets[i + 1])] for i in range(self.n_lists)]


def plan(overlays, image_width, image_height, layout="rgb", out=None, n_frames=None):
    """``overlays``: one overlay dict (shared by every frame of the call, as in the reference), or a
    list with one dict per frame (frames of several streams batched together).  Identical dicts
    share one draw list.  ``out``: optional uint8 ndarray to plan into (e.g. over pinned memory).
    ``n_frames``: frames of the call when one dict is shared (default 1).  -> Plan"""
    shared = isinstance(overlays, dict)
    dicts = [overlays] if shared else list(overlays)
    lists, list_of_frame, seen = [], [], {}
    for o in dicts:
        key = id(o)
        li = seen.get(key)
        if li is None:
            li = seen[key] = len(lists)
            lists.append(o)
        list_of_frame.append(li)
    if shared and n_frames:
        list_of_frame = [0] * int(n_frames)
    rects, labels, offsets = [], [], [0]
    for o in lists:
        offsets.append(offsets[-1] + _objects(o, rects, labels))
    n_obj = offsets[-1]
    n_lists, n_frames = len(lists), len(list_of_frame)
    # every character is at most one op, every object one more
    text = "\0".join(labels)
    try:
        raw = text.encode("ascii")
    except UnicodeEncodeError:           # cv2 sees the UTF-8 bytes, one glyph each ('?' outside 32..126)
        raw = b"\0".join(_encode(t) for t in labels)
    cap = len(raw) + 1
    ops_bytes = max(cap, 1) * 32
    need = ops_bytes + 4 * (n_lists + 1) + 4 * n_frames
    if out is None or out.nbytes < need:
        out = np.empty(need, np.uint8)
    base = out.ctypes.data
    c_offsets = array("i", offsets)
    c_rects = array("i", rects or (0,))
    info = OverlayPlanInfo()
    check(_native.lib().aiko_b200_overlay_plan_host(
        n_lists, c_offsets.buffer_info()[0], c_rects.buffer_info()[0], raw + b"\0", int(image_width), int(image_height),
        LAYOUTS[layout], base, cap, base + ops_bytes, ctypes.byref(info)))
    out[ops_bytes + 4 * (n_lists + 1):need].view(np.int32)[:] = list_of_frame
    p = Plan()
    p.buffer, p.n_ops, p.n_lists, p.n_frames, p.info = out[:need], info.n_ops, n_lists, n_frames, info
    p.width, p.height, p.ops_bytes = int(image_width), int(image_height), ops_bytes
    return p


def build_ops(overlay, image_width, image_height, layout="rgb", atlas=None):
    """-> list[OverlayOp] for one frame of size image_width x image_height."""
    return plan(overlay, image_width, image_height, layout).ops()[0]


def label_origin(atlas, text, x, y, w, h, image_width):
    """image_io.py:140-150 -> (tx, ty), the bottom-left text origin given to cv2.putText (the rule
    the library applies; kept here as its executable statement for the tests)."""
    tw, th = (atlas or TextAtlas.shared()).text_size(text)
    tx = x
    if y > th * 1.5:
        ty = int(y - th / 2 - 2)
    else:
        ty = y + h + th + 2
    if x + tw > image_width:
        tx = x - tw
        ty = y + th
    return tx, ty


def gather_objects(overlays):
    """-> (n_lists, offsets array('i'), rects array('i'), labels bytes, list_of_frame or None) for one
    overlay dict (shared by all frames: list_of_frame None) or a list of dicts (one per frame)."""
    if isinstance(overlays, dict):
        lists, list_of_frame = [overlays], None
    else:
        lists, list_of_frame, seen = [], [], {}
        for o in overlays:
            li = seen.get(id(o))
            if li is None:
                li = seen[id(o)] = len(lists)
                lists.append(o)
            list_of_frame.append(li)
    rects, labels, offsets = [], [], [0]
    for o in lists:
        offsets.append(offsets[-1] + _objects(o, rects, labels))
    text = "\0".join(labels)
    try:
        raw = text.encode("ascii")
    except UnicodeEncodeError:
        raw = b"\0".join(_encode(t) for t in labels)
    return len(lists), array("i", offsets), array("i", rects or (0,)), raw + b"\0", list_of_frame
