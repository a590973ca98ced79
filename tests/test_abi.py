"""CPU: the C-ABI library loads and exports every symbol include/aiko_b200.h declares."""
import ctypes
import os
import re

from aiko_services_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "aiko_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(aiko_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_bindings_agree():
    assert declared_symbols() == sorted(_native.SIGNATURES)


def test_library_exports_every_declared_symbol():
    handle = ctypes.CDLL(_native.LIB_PATH)
    for name in declared_symbols():
        assert getattr(handle, name) is not None


def test_version_and_error_string_without_gpu():
    lib = _native.lib()
    assert lib.aiko_b200_abi_version() == 2
    # argument validation happens before any CUDA call, so it is testable on CPU
    rc = lib.aiko_b200_swap_rb_u8(None, None, -1, None)
    assert rc == _native.ERR_INVALID_ARGUMENT
    assert b"n_pixels" in lib.aiko_b200_last_error()
    rc = lib.aiko_b200_resize_linear_u8(1, 1, 4, 4, 2, 8, 32, 1, 2, 2, 4, 8, 0, None, None)
    assert rc == _native.ERR_INVALID_ARGUMENT and b"channels" in lib.aiko_b200_last_error()
    assert lib.aiko_b200_resize_linear_u8(None, 0, 4, 4, 3, 12, 48, None, 2, 2, 6, 12, 0, None, None) == 0


def test_overlay_op_struct_layout():
    assert ctypes.sizeof(_native.OverlayOp) == 32
    assert ctypes.sizeof(_native.Overlay) == 96    # 6 pointers + 10 int32 + uint64


def test_zero_copy_eligibility_rule():
    """Frames may be read from pinned host memory by the kernel only when every needed source row is
    fetched once with bulk row copies: RGB, 16-byte aligned rows, scale x <= 4, scale y >= 2."""
    from aiko_services_b200 import ops
    assert ops.reads_source_rows_once((1080, 1920, 3), 640, 480)
    assert not ops.reads_source_rows_once((1080, 1920, 3), 1280, 720)      # rows would be fetched twice
    assert not ops.reads_source_rows_once((1080, 1918, 3), 640, 480)       # unaligned rows: gathers
    assert not ops.reads_source_rows_once((1080, 1920), 640, 480)          # gray: generic kernel
    assert not ops.reads_source_rows_once((2160, 3840, 3), 224, 224)       # scale x > 4: gathers


def test_jpeg_header_parser_survives_hostile_input():
    """aiko_b200_jpeg_gpu_info_host on damaged and truncated headers: an error code or the header's numbers,
    never a crash (host code only -- runs without a GPU)."""
    import ctypes
    import io

    import numpy as np
    from PIL import Image
    from aiko_services_b200 import _native
    lib = _native.lib()
    rng = np.random.default_rng(5)
    frame = rng.integers(0, 256, (40, 56, 3), dtype=np.uint8)
    blobs = []
    for kw in ({}, {"restart_marker_rows": 1}, {"progressive": True}, {"subsampling": 0}):
        buf = io.BytesIO()
        Image.fromarray(frame, "RGB").save(buf, "JPEG", quality=80, **kw)
        blobs.append(buf.getvalue())
    vals = [ctypes.c_int() for _ in range(6)]
    results = set()
    for trial in range(3000):
        blob = bytearray(blobs[trial % len(blobs)])
        head = min(len(blob), 700)
        for _ in range(int(rng.integers(1, 8))):
            blob[int(rng.integers(0, head))] = int(rng.integers(256))
        if trial % 3 == 0:
            blob = blob[:int(rng.integers(0, head))]
        raw = bytes(blob)
        rc = lib.aiko_b200_jpeg_gpu_info_host(raw, len(raw), *[ctypes.byref(v) for v in vals])
        results.add(rc)
        assert rc in (_native.OK, _native.ERR_INVALID_ARGUMENT, _native.ERR_UNSUPPORTED)
    assert _native.OK in results and _native.ERR_INVALID_ARGUMENT in results


def test_jpeg_header_parser_refuses_rgb_coded_files():
    """A JPEG whose three components are RGB (Adobe marker, colour transform 0) is not YCbCr: the own decoder
    must refuse it (the caller keeps its other decoder) rather than decode wrong colours."""
    import ctypes
    import io

    import numpy as np
    from PIL import Image
    from aiko_services_b200 import _native
    frame = np.random.default_rng(0).integers(0, 256, (40, 56, 3), dtype=np.uint8)
    vals = [ctypes.c_int() for _ in range(6)]
    buf = io.BytesIO()
    try:
        Image.fromarray(frame, "RGB").save(buf, "JPEG", keep_rgb=True)
    except TypeError:                       # a Pillow without keep_rgb
        return
    raw = buf.getvalue()
    assert _native.lib().aiko_b200_jpeg_gpu_info_host(raw, len(raw), *[ctypes.byref(v) for v in vals]) == _native.ERR_UNSUPPORTED
    buf = io.BytesIO()                      # comment, EXIF and ICC segments are skipped
    Image.fromarray(frame, "RGB").save(buf, "JPEG", comment=b"hello", icc_profile=b"x" * 3000,
                                       exif=b"Exif\x00\x00II*\x00\x08\x00\x00\x00\x00\x00")
    raw = buf.getvalue()
    assert _native.lib().aiko_b200_jpeg_gpu_info_host(raw, len(raw), *[ctypes.byref(v) for v in vals]) == _native.OK
    assert [v.value for v in vals[:5]] == [56, 40, 3, 2, 2]
