"""Generate tests/golden/hershey_atlas_cv2.npz with OpenCV itself (run where cv2 is installed).

For every printable ASCII character and each of the four sub-pixel pen phases that
cv2.putText(FONT_HERSHEY_SIMPLEX, 0.75, (0, 255, 255), 2, LINE_AA) can reach, the character is
rendered by cv2 on all 256 constant backgrounds; per pixel that gives the exact transfer
function of the drawing (old value -> new value) for colour components 0 and 255.  The file is
the oracle for the library's own rasteriser (tests/test_text_raster.py): patch p of the library
must realise the same per-pixel transfer functions.

    python tests/golden/make_atlas_golden.py
"""
import os

import numpy as np

FONT_SCALE, THICKNESS = 0.75, 2
COMPONENTS = (0, 255)
FIRST_CHAR, LAST_CHAR = 32, 126
PAD_X, PAD_Y = 8, 40
PATCH_W, PATCH_H = 40, 48
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hershey_atlas_cv2.npz")


def _phase_prefixes(advances):
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
    out = np.empty((256, height, width, 3), np.uint8)
    for b in range(256):
        img = np.full((height, width, 3), b, np.uint8)
        cv2.putText(img, text, org, cv2.FONT_HERSHEY_SIMPLEX, FONT_SCALE, colour_bgr, THICKNESS, cv2.LINE_AA)
        out[b] = img
    return out


def main(path=OUT):
    import cv2
    font = cv2.FONT_HERSHEY_SIMPLEX
    n_chars = LAST_CHAR - FIRST_CHAR + 1
    advances = np.array([cv2.getTextSize(chr(FIRST_CHAR + i), font, 1.0, 0)[0][0] for i in range(n_chars)], np.int32)
    text_height = cv2.getTextSize("A", font, FONT_SCALE, THICKNESS)[0][1]
    prefixes = _phase_prefixes(advances)
    identity = np.arange(256, dtype=np.uint8)
    lut_index = {identity.tobytes(): 0}
    luts = [identity]
    n_patches = 1 + n_chars * 4 * len(COMPONENTS)
    patch_ids = np.zeros((n_patches, PATCH_H, PATCH_W), np.uint16)
    org_x, org_y = 16, PAD_Y
    for ci in range(n_chars):
        for q in range(4):
            prefix = prefixes[q]
            units = sum(int(advances[ord(c) - FIRST_CHAR]) for c in prefix)
            pen_q = 3 * units
            assert pen_q % 4 == q
            x0 = org_x + pen_q // 4 - PAD_X
            width = x0 + PATCH_W + 16
            tr = _render_transfer(cv2, prefix + chr(FIRST_CHAR + ci), (org_x, org_y), width, PATCH_H + 8, (0, 255, 255))
            region = tr[:, :PATCH_H, x0:x0 + PATCH_W, :]
            for k, comp in enumerate(COMPONENTS):
                chan = 0 if comp == 0 else 1
                pidx = 1 + (ci * 4 + q) * len(COMPONENTS) + k
                fn = region[:, :, :, chan]
                ys, xs = np.nonzero((fn != identity[:, None, None]).any(0))
                for y, x in zip(ys, xs):
                    key = fn[:, y, x].tobytes()
                    idx = lut_index.get(key)
                    if idx is None:
                        idx = len(luts)
                        lut_index[key] = idx
                        luts.append(fn[:, y, x].copy())
                    patch_ids[pidx, y, x] = idx
    np.savez_compressed(path, advances=advances, text_height=np.int32(text_height), patch_ids=patch_ids,
                        luts=np.stack(luts).astype(np.uint8), cv2_version=np.array(cv2.__version__))
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
