"""Diagnostic: rounds the boundary-state iteration of jpeg_sync_huffman_kernel needs (AIKO_B200_JPEG_DEBUG)."""
import os
import sys

import torch

os.environ["AIKO_B200_JPEG_DEBUG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jpeg_decode import blob, pinned  # noqa: E402
from aiko_services_b200 import jpeg  # noqa: E402

dev = torch.device("cuda:0")
for kw in ({}, {"restart_marker_rows": 1}, {"restart_marker_blocks": 30}):
    blobs = [pinned(blob(500 + k, **kw)) for k in range(32)]
    print(kw, file=sys.stderr)
    jpeg.GpuJpegDecoder(dev).decode(blobs, check=True)
