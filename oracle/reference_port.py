"""The reference elements' process_frame bodies as plain functions -- TEST INFRASTRUCTURE.

Same third-party calls, in the same order, as R/elements/media/image_io.py (cv2 / Pillow do
the arithmetic there too).  Used by tests as a second check next to the C oracle, and by
bench.py as the timed CPU reference arm on the GPU box, where /root/reference is absent but
cv2 / Pillow (the libraries the reference delegates to) are installed.
"""
import numpy as np


def image_resize(images, resolution):
    """ImageResize.process_frame, image_io.py:208-220"""
    import cv2
    width, height = resolution.split("x")
    out = []
    for image in images:
        if isinstance(image, np.ndarray):
            out.append(cv2.resize(image, (int(width), int(height))))
        else:
            out.append(image.resize((int(width), int(height))))
    return out


def image_overlay(images, overlay, color=(0, 255, 255), font_scale=0.75, thickness=2, threshold=0.0):
    """ImageOverlay.process_frame, image_io.py:107-163"""
    import cv2
    font = cv2.FONT_HERSHEY_SIMPLEX
    result = []
    for image in images:
        image_width = image.shape[1]
        gray = image.ndim == 2
        bgr = cv2.cvtColor(image, cv2.COLOR_GRAY2BGR if gray else cv2.COLOR_RGB2BGR)
        if "rectangles" in overlay:
            rectangles = overlay["rectangles"]
            objects = overlay["objects"] if "objects" in overlay else [{}] * len(rectangles)
            for obj, rect in zip(objects, rectangles):
                name = obj.get("name", None)
                confidence = obj.get("confidence", 1.0)
                if confidence > threshold:
                    x, y, w, h = int(rect["x"]), int(rect["y"]), int(rect["w"]), int(rect["h"])
                    if name:
                        text = f"{name}: {confidence:0.2f}"
                        (tw, th), _ = cv2.getTextSize(text, font, font_scale, thickness)
                        tx = x
                        ty = int(y - th / 2 - 2) if y > th * 1.5 else y + h + th + 2
                        if x + tw > image_width:
                            tx, ty = x - tw, y + th
                        bgr = cv2.putText(bgr, text, (tx, ty), font, font_scale, color, thickness, cv2.LINE_AA)
                    bgr = cv2.rectangle(bgr, (x, y), (x + w, y + h), color, thickness)
        result.append(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY if gray else cv2.COLOR_BGR2RGB))
    return result


def chain_resize_overlay_bgr(images, resolution, overlay):
    """BASELINE config 2: ImageResize -> ImageOverlay -> RGB2BGR for the encoder
    (R/elements/media/video_io.py:299)."""
    import cv2
    frames = image_overlay(image_resize(images, resolution), overlay)
    return [cv2.cvtColor(f, cv2.COLOR_RGB2BGR) for f in frames]
