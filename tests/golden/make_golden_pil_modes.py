"""Generate tests/golden/golden_pil_modes.npz: the UNMODIFIED reference's ImageResize (PIL branch,
R/elements/media/image_io.py:218) on PIL images of modes RGBA, LA, L, CMYK, P and 1.  Runs only where
/root/reference exists.

    python tests/golden/make_golden_pil_modes.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, HERE)

import cases  # noqa: E402
from oracle import ref_harness  # noqa: E402

IMAGE_IO = "aiko_services.elements.media.image_io"


def main():
    import PIL
    store = {"versions": np.array(f"Pillow {PIL.__version__}")}
    for name, seed, mode, sh, sw, dw, dh in cases.RESIZE_PIL_MODES:
        img = cases.pil_image(seed, mode, sh, sw)
        out = ref_harness.run_reference_pipeline(
            [("ImageResize", IMAGE_IO, ["images"], ["images"], {"resolution": f"{dw}x{dh}"}),
             ("ImageOutput", IMAGE_IO, ["images"], ["images"], None)], [{"images": [img]}])
        result = out[0]["images"][0]
        assert result.mode == mode, (name, result.mode)
        arr = np.asarray(result)
        store[name] = arr.astype(np.uint8) * 255 if mode == "1" else arr
    path = os.path.join(HERE, "golden_pil_modes.npz")
    np.savez_compressed(path, **store)
    print(path, os.path.getsize(path), "bytes;", store["versions"])


if __name__ == "__main__":
    main()
