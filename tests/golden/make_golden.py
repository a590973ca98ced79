"""Generate tests/golden/golden.npz by running the UNMODIFIED reference elements.

Runs only where /root/reference exists (the build container).  The reference's own
PipelineImpl drives its own ImageResize / ImageOverlay on the seeded frames of cases.py;
small outputs are stored whole, large ones as SHA-256 plus a crop.  Library versions are
recorded: the arithmetic lives in cv2 / Pillow, which the reference does not vendor.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, HERE)

import cases  # noqa: E402
from oracle import ref_harness  # noqa: E402

IMAGE_IO = "aiko_services.elements.media.image_io"
BIG = 200_000  # bytes: larger outputs are stored as hash + crop


def _store(store, name, arr):
    arr = np.ascontiguousarray(arr)
    store[name + "/sha256"] = np.array(hashlib.sha256(arr.tobytes()).hexdigest())
    store[name + "/shape"] = np.array(arr.shape)
    if arr.nbytes <= BIG:
        store[name + "/full"] = arr
    else:
        store[name + "/crop"] = arr[:32, :48].copy()
        store[name + "/crop_end"] = arr[-16:, -24:].copy()


def main():
    import cv2, PIL
    from PIL import Image
    store = {"versions": np.array(f"cv2 {cv2.__version__}; Pillow {PIL.__version__}; numpy {np.__version__}")}

    for name, seed, sh, sw, c, dw, dh in cases.RESIZE_NDARRAY:
        img = cases.frame(seed, sh, sw, c)
        out = ref_harness.run_reference_pipeline(
            [("ImageResize", IMAGE_IO, ["images"], ["images"], {"resolution": f"{dw}x{dh}"}),
             ("ImageOutput", IMAGE_IO, ["images"], ["images"], None)], [{"images": [img]}])
        _store(store, "resize/" + name, out[0]["images"][0])

    for name, seed, sh, sw, dw, dh in cases.RESIZE_PIL:
        img = Image.fromarray(cases.frame(seed, sh, sw, 3), "RGB")
        out = ref_harness.run_reference_pipeline(
            [("ImageResize", IMAGE_IO, ["images"], ["images"], {"resolution": f"{dw}x{dh}"}),
             ("ImageOutput", IMAGE_IO, ["images"], ["images"], None)], [{"images": [img]}])
        _store(store, "resize_pil/" + name, np.asarray(out[0]["images"][0]))

    for name, seed, h, w, c, overlay in cases.OVERLAY:
        img = cases.frame(seed, h, w, c)
        out = ref_harness.run_reference_pipeline(
            [("ImageOverlay", IMAGE_IO, ["images", "overlay"], ["images"], None),
             ("ImageOutput", IMAGE_IO, ["images"], ["images"], None)], [{"images": [img], "overlay": overlay}])
        _store(store, "overlay/" + name, out[0]["images"][0])

    # the BASELINE config-2 chain: ImageResize(1080p -> 640x480) -> ImageOverlay, 2 frames per batch
    imgs = [cases.frame(2000 + i, 1080, 1920, 3) for i in range(2)]
    out = ref_harness.run_reference_pipeline(
        [("ImageResize", IMAGE_IO, ["images"], ["images"], {"resolution": "640x480"}),
         ("ImageOverlay", IMAGE_IO, ["images", "overlay"], ["images"], None),
         ("ImageOutput", IMAGE_IO, ["images"], ["images"], None)],
        [{"images": imgs, "overlay": cases.OVERLAY_4}])
    for i, o in enumerate(out[0]["images"]):
        _store(store, f"chain/cfg2_{i}", o)

    path = os.path.join(HERE, "golden.npz")
    np.savez_compressed(path, **store)
    print(path, os.path.getsize(path), "bytes;", store["versions"])


if __name__ == "__main__":
    main()
