"""Host-side translation of ImageOverlay's per-object logic into a GPU draw list.

Follows R/elements/media/image_io.py:119-155 step by step: objects and rectangles are
zipped, objects at or below the confidence threshold are skipped, the label text is
``f"{name}: {confidence:0.2f}"``, its position follows the reference's placement rule, and
for each object the label is drawn before its rectangle.  The result is a list of
``OverlayOp`` records (include/aiko_b200.h) that the kernels replay in order.
"""
from ._native import OverlayOp
from .text_atlas import COMPONENTS, TextAtlas

OP_RECT, OP_GLYPH = 0, 1

COLOR_BGR = (0, 255, 255)   # image_io.py:101, drawn on the BGR copy of the frame
THRESHOLD = 0.0             # image_io.py:105

# memory-channel colour of the stored frame for each layout the kernels produce
CHANNEL_COLOURS = {
    "rgb": (COLOR_BGR[2], COLOR_BGR[1], COLOR_BGR[0]),   # frame stays RGB (image_io.py:117,160)
    "bgr": COLOR_BGR,                                    # frame stored swapped, or the gray round trip
}


def label_origin(atlas, text, x, y, w, h, image_width):
    """image_io.py:140-150 -> (tx, ty), the bottom-left text origin given to cv2.putText."""
    tw, th = atlas.text_size(text)
    tx = x
    if y > th * 1.5:
        ty = int(y - th / 2 - 2)
    else:
        ty = y + h + th + 2
    if x + tw > image_width:
        tx = x - tw
        ty = y + th
    return tx, ty


def build_ops(overlay, image_width, image_height, layout="rgb", atlas=None):
    """-> list[OverlayOp] for one frame of size image_width x image_height.

    ``layout`` names the channel order of the frame the ops will be applied to.
    """
    atlas = atlas or TextAtlas.shared()
    colour = CHANNEL_COLOURS[layout]
    for comp in colour:
        if comp not in COMPONENTS:
            raise ValueError(f"glyph atlas has no tables for colour component {comp}")
    ops = []
    if "rectangles" not in overlay:
        return ops
    rectangles = overlay["rectangles"]
    objects = overlay["objects"] if "objects" in overlay else [{}] * len(rectangles)
    for obj, rectangle in zip(objects, rectangles):
        name = obj.get("name", None)
        confidence = obj.get("confidence", 1.0)
        if not confidence > THRESHOLD:
            continue
        x, y, w, h = int(rectangle["x"]), int(rectangle["y"]), int(rectangle["w"]), int(rectangle["h"])
        if name:
            text = f"{name}: {confidence:0.2f}"
            tx, ty = label_origin(atlas, text, x, y, w, h, image_width)
            for ch, phase, px, py in atlas.glyphs(text, tx, ty):
                pidx = [atlas.patch_index(ch, phase, comp) for comp in colour]
                keys = [atlas.border_key(p, px, py, image_width, image_height) for p in set(pidx)]
                key = next((k for k in keys if k is not None), None)
                if key is not None:
                    pidx = atlas.border_variant(ch, phase, colour, key)
                boxes = [atlas.bbox[p] for p in pidx]
                boxes = [b for b in boxes if b[2] >= b[0]]
                if not boxes:
                    continue
                bx0 = min(b[0] for b in boxes); by0 = min(b[1] for b in boxes)
                bx1 = max(b[2] for b in boxes); by1 = max(b[3] for b in boxes)
                ax0, ay0, ax1, ay1 = px + bx0, py + by0, px + bx1, py + by1
                # clip to the image; skip glyphs that fall entirely outside
                cx0, cy0 = max(ax0, 0), max(ay0, 0)
                cx1, cy1 = min(ax1, image_width - 1), min(ay1, image_height - 1)
                if cx1 < cx0 or cy1 < cy0:
                    continue
                ops.append(OverlayOp(OP_GLYPH, int(cx0), int(cy0), int(cx1), int(cy1),
                                     int(pidx[0]) | (int(pidx[1]) << 16), int(pidx[2]),
                                     int(cx0 - px) | (int(cy0 - py) << 16)))
        x0, y0, x1, y1 = x, y, x + w, y + h
        xa, xb = min(x0, x1), max(x0, x1)
        ya, yb = min(y0, y1), max(y0, y1)
        if xb + 1 < 0 or yb + 1 < 0 or xa - 1 >= image_width or ya - 1 >= image_height:
            continue
        # keep coordinates in a range where +-2 arithmetic cannot overflow int32 on the device
        lim = 1 << 24
        xa, ya, xb, yb = (max(-lim, min(lim, v)) for v in (xa, ya, xb, yb))
        ops.append(OverlayOp(OP_RECT, xa, ya, xb, yb,
                             colour[0] | (colour[1] << 8) | (colour[2] << 16), 0, 0))
    return ops
