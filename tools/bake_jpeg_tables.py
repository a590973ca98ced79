"""Writes the standard JPEG tables (ITU-T T.81 Annex K: the two example quantisation tables and the four
typical Huffman tables that libjpeg installs by default) as a C include file.  They are read back out of a
file written by Pillow at quality 50 (scale factor 100 %: the quantisation tables are then Annex K's own).

    python tools/bake_jpeg_tables.py aiko_services_b200/csrc/jpeg_std_tables.inc oracle/jpeg_std_tables.inc
"""
import io
import sys

import numpy as np
from PIL import Image

ZIGZAG = [0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21,
          28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61,
          54, 47, 55, 62, 63]


def main():
    buf = io.BytesIO()
    Image.fromarray(np.zeros((16, 16, 3), np.uint8), "RGB").save(buf, "JPEG", quality=50)
    b = buf.getvalue()
    pos, quant, huff = 2, {}, {}
    while pos < len(b):
        m, length = b[pos + 1], (b[pos + 2] << 8) | b[pos + 3]
        seg = b[pos + 4:pos + 2 + length]
        if m == 0xDB:
            natural = [0] * 64
            for k in range(64):
                natural[ZIGZAG[k]] = seg[1 + k]
            quant[seg[0] & 15] = natural
        elif m == 0xC4:
            counts = list(seg[1:17])
            huff[seg[0]] = (counts, list(seg[17:17 + sum(counts)]))
        elif m == 0xDA:
            break
        pos += 2 + length
    assert quant[0][:4] == [16, 11, 10, 16] and quant[1][:4] == [17, 18, 24, 47], "not the Annex K tables"
    lines = ["// Standard JPEG tables, ITU-T T.81 Annex K (written by tools/bake_jpeg_tables.py).",
             "// Quantisation tables in natural (row-major) order; Huffman tables as BITS[1..16] + HUFFVAL."]
    for name, t in (("kStdLuminanceQuant", quant[0]), ("kStdChrominanceQuant", quant[1])):
        lines.append(f"static const unsigned char {name}[64] = {{{', '.join(map(str, t))}}};")
    for name, key in (("DcLuminance", 0x00), ("AcLuminance", 0x10), ("DcChrominance", 0x01), ("AcChrominance", 0x11)):
        counts, vals = huff[key]
        lines.append(f"static const unsigned char kStd{name}Bits[16] = {{{', '.join(map(str, counts))}}};")
        lines.append(f"static const unsigned char kStd{name}Vals[{len(vals)}] = {{{', '.join(map(str, vals))}}};")
    text = "\n".join(lines) + "\n"
    for path in sys.argv[1:]:
        open(path, "w").write(text)


if __name__ == "__main__":
    main()
