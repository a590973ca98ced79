"""GPU: the allocation-free bicubic resize entry point (aiko_b200_resize_bicubic_aa_u8_noalloc).

It exists for callers that cannot allocate device memory at the call site -- for example an
element whose process_frame runs on the pipeline's single event-loop thread, where a cudaMalloc
would stall every stream -- so its weight tables sit in an array built into the library.  This
file is named to be collected last: if it fails, a run that stops at the first failure has
already run every other test.

Checked through the C ABI: two calls in a row (the second with smaller tables, so whatever
the first left behind is still in the array) each give exactly the bytes of the existing
workspace entry point, and a size pair whose tables exceed the built-in array is refused.
"""
import ctypes

import numpy as np
import pytest
import torch

from aiko_services_b200 import _native, ops
from golden import cases

pytestmark = pytest.mark.gpu


def _noalloc(lib, src, dst_h, dst_w):
    n, h, w, c = src.shape
    dst = torch.empty((n, dst_h, dst_w, c), dtype=torch.uint8, device=src.device)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.aiko_b200_resize_bicubic_aa_u8_noalloc(src.data_ptr(), n, h, w, c, dst.data_ptr(), dst_h, dst_w, stream)
    return rc, dst


def test_noalloc_matches_workspace_entry_point_twice_in_a_row(cuda):
    lib = _native.lib()
    shapes = [((3, 1080, 1920, 3), (480, 640)),      # 237 KB of tables (coefficients + tensor-core operand tables)
              ((2, 97, 131, 3), (41, 50)),           # much smaller tables: the rest of the array keeps call 1's
              ((2, 60, 64, 1), (20, 64))]            # width unchanged: identity horizontal table
    for i, (shape, (dh, dw)) in enumerate(shapes):
        n, h, w, c = shape
        frames = np.stack([cases.frame(9100 + 10 * i + k, h, w, c).reshape(h, w, c) for k in range(n)])
        src = torch.from_numpy(frames).to(cuda)
        rc, got = _noalloc(lib, src, dh, dw)
        assert rc == 0, lib.aiko_b200_last_error()
        want = ops.resize_bicubic_aa(src, dw, dh)    # existing entry point, caller-owned workspace
        torch.cuda.synchronize()
        assert torch.equal(got, want)                # bit-exact, the bar of every image test here


def test_noalloc_refuses_sizes_beyond_the_built_in_scratch(cuda):
    lib = _native.lib()
    assert lib.aiko_b200_resize_bicubic_workspace_bytes(8000, 8000, 4000, 4000) > 256 * 1024
    src = torch.zeros((1, 4, 4, 3), dtype=torch.uint8, device=cuda)   # never read: the size check comes first
    dst = torch.zeros((1, 4, 4, 3), dtype=torch.uint8, device=cuda)
    rc = lib.aiko_b200_resize_bicubic_aa_u8_noalloc(src.data_ptr(), 1, 8000, 8000, 3, dst.data_ptr(), 4000, 4000, None)
    assert rc == _native.ERR_UNSUPPORTED
    message = lib.aiko_b200_last_error().decode()
    assert "built-in scratch" in message and "workspace" in message
    assert lib.aiko_b200_resize_bicubic_aa_u8_noalloc(None, -1, 8, 8, 3, None, 4, 4, None) == _native.ERR_INVALID_ARGUMENT
