"""Loader for libaiko_b200.so -- the C-ABI boundary (include/aiko_b200.h).

There is no CPU fallback: if the library is missing, or no B200 is visible when a
compute entry point is called, the product path raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaiko_b200.so")
CSRC = os.path.join(_HERE, "csrc")

OK = 0
ERR_INVALID_ARGUMENT = -1
ERR_UNSUPPORTED = -2
ERR_CUDA = -3
ERR_WORKSPACE = -4
ATLAS_GREW = 1          # aiko_b200_resize_linear_objects_u8: planned, not drawn


class NativeError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libaiko_b200 error {code}: {message}")
        self.code = code


class OverlayOp(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int32), ("x0", ctypes.c_int32), ("y0", ctypes.c_int32),
                ("x1", ctypes.c_int32), ("y1", ctypes.c_int32), ("p0", ctypes.c_uint32),
                ("p1", ctypes.c_uint32), ("p2", ctypes.c_uint32)]


class Overlay(ctypes.Structure):
    _fields_ = [("ops", ctypes.c_void_p), ("op_offsets", ctypes.c_void_p), ("list_of_frame", ctypes.c_void_p),
                ("patch_ids", ctypes.c_void_p), ("luts", ctypes.c_void_p),
                ("patch_w", ctypes.c_int32), ("patch_h", ctypes.c_int32), ("n_patches", ctypes.c_int32),
                ("n_lists", ctypes.c_int32), ("n_frames", ctypes.c_int32), ("max_ops_per_list", ctypes.c_int32),
                ("max_marks_per_list", ctypes.c_int32), ("max_cells_per_list", ctypes.c_int32),
                ("frame_width", ctypes.c_int32), ("frame_height", ctypes.c_int32),
                ("upload_from_host", ctypes.c_void_p), ("upload_bytes", ctypes.c_uint64)]


class OverlayPlanInfo(ctypes.Structure):
    _fields_ = [("n_ops", ctypes.c_int32), ("max_ops_per_list", ctypes.c_int32),
                ("max_marks_per_list", ctypes.c_int32), ("max_cells_per_list", ctypes.c_int32),
                ("atlas_version", ctypes.c_uint64), ("atlas_epoch", ctypes.c_uint64)]


class AtlasView(ctypes.Structure):
    _fields_ = [("patch_ids_host", ctypes.c_void_p), ("luts_host", ctypes.c_void_p),
                ("patch_w", ctypes.c_int32), ("patch_h", ctypes.c_int32), ("n_patches", ctypes.c_int32),
                ("n_luts", ctypes.c_int32), ("max_patches", ctypes.c_int32), ("max_luts", ctypes.c_int32),
                ("version", ctypes.c_uint64), ("epoch", ctypes.c_uint64)]


_vp, _i, _i64, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t
_ovp = ctypes.POINTER(Overlay)

# name -> argtypes.  Every symbol declared in include/aiko_b200.h must be listed here;
# tests/test_abi.py checks the two against each other.
SIGNATURES = {
    "aiko_b200_abi_version": [],
    "aiko_b200_last_error": [],
    "aiko_b200_device_info": [ctypes.c_char_p, _sz, ctypes.POINTER(_i), ctypes.POINTER(_i)],
    "aiko_b200_overlay_plan_host": [_i, _vp, _vp, ctypes.c_char_p, _i, _i, _i, _vp, _i, _vp,
                                    ctypes.POINTER(OverlayPlanInfo)],
    "aiko_b200_overlay_plan_struct_host": [_i, _vp, _vp, ctypes.c_char_p, _i, _i, _i, _i, _vp, _vp, _sz, _vp, _ovp,
                                           ctypes.POINTER(OverlayPlanInfo)],
    "aiko_b200_resize_linear_objects_u8": [_vp, _i, _i, _i, _i, _i64, _i64, _vp, _i, _i, _i64, _i64, _i, _i, _vp, _vp,
                                           ctypes.c_char_p, _i, _vp, _vp, _sz, _vp, _ovp,
                                           ctypes.POINTER(OverlayPlanInfo), ctypes.c_uint64, ctypes.c_uint64, _vp],
    "aiko_b200_overlay_upload": [_vp, _vp, _sz, _vp],
    "aiko_b200_overlay_text_size_host": [ctypes.c_char_p, ctypes.POINTER(_i), ctypes.POINTER(_i)],
    "aiko_b200_overlay_cells_host": [_vp, _i, _i, _i, ctypes.POINTER(_i), ctypes.POINTER(ctypes.c_longlong)],
    "aiko_b200_overlay_marks_host": [_vp, _i, _i, _vp],
    "aiko_b200_overlay_atlas_host": [ctypes.POINTER(AtlasView)],
    "aiko_b200_overlay_glyph_metrics_host": [_vp, _vp, _vp],
    "aiko_b200_overlay_boxes_workspace_bytes": [_i, _i, _i],
    "aiko_b200_overlay_boxes_u8": [_vp, _i, _i, _i, _i64, _i, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _i,
                                   _vp, _sz, _vp, _vp],
    "aiko_b200_overlay_variants_host": [_vp, _vp, _vp, _i, _vp, _vp],
    "aiko_b200_resize_linear_u8": [_vp, _i, _i, _i, _i, _i64, _i64, _vp, _i, _i, _i64, _i64, _i, _ovp, _vp],
    "aiko_b200_resize_linear_rows_host": [_i, _i, _vp],
    "aiko_b200_copy_rows_h2d": [_vp, _vp, _i, _i64, _i64, _vp, _i, _vp],
    "aiko_b200_resize_bicubic_workspace_bytes": [_i, _i, _i, _i],
    "aiko_b200_resize_bicubic_prepare": [_vp, _sz, _i, _i, _i, _i, _vp],
    "aiko_b200_resize_bicubic_aa_u8": [_vp, _i, _i, _i, _i, _vp, _i, _i, _vp, _sz, _vp],
    "aiko_b200_resize_bicubic_aa_u8_noalloc": [_vp, _i, _i, _i, _i, _vp, _i, _i, _vp],
    "aiko_b200_premultiply_u8": [_vp, _vp, _i64, _i, _i, _vp],
    "aiko_b200_gather_nearest_u8": [_vp, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _vp],
    "aiko_b200_resize_norm_f32_nchw": [_vp, _i, _i, _i, _vp, _i, _i, ctypes.POINTER(ctypes.c_float),
                                       ctypes.POINTER(ctypes.c_float), _vp],
    "aiko_b200_audio_workspace_bytes": [_i],
    "aiko_b200_audio_prepare": [_vp, _sz, _i, _vp],
    "aiko_b200_audio_chain": [_vp, _i, _i, _i, ctypes.c_double, _vp, _vp, ctypes.POINTER(_i), _vp, _vp, _vp, _i, _i,
                              _vp, _sz, _vp],
    "aiko_b200_audio_filter_f64": [_vp, _vp, _i, _i, _i64, _vp, _vp, _vp, _vp],
    "aiko_b200_audio_bands_f64": [_vp, _vp, _i, _i, _i64, _vp, _vp, _vp],
    "aiko_b200_jpeg_gpu_info_host": [ctypes.c_char_p, _sz] + [ctypes.POINTER(_i)] * 6,
    "aiko_b200_jpeg_gpu_staging_bytes": [_i],
    "aiko_b200_jpeg_gpu_workspace_bytes": [_i, _i, _i, _i, _i, _i, _sz, _i],
    "aiko_b200_jpeg_gpu_decode_rgb_u8": [_vp, _vp, _i, _vp, _i, _i, _i64, _vp, _sz, _vp, _sz, _sz, _i, _vp],
    "aiko_b200_jpeg_gpu_status": [_vp, _i, _i, _i, _i, _i, _i, _sz, _i, _vp],
    "aiko_b200_jpeg_gpu_encode_workspace_bytes": [_i, _i, _i, _i, _i, _i, _sz],
    "aiko_b200_jpeg_gpu_encode_u8": [_vp, _i, _i, _i, _i, _i64, _i, _i, _i, _vp, _sz, _sz, _vp, _sz, _vp, _vp],
    "aiko_b200_overlay_u8": [_vp, _vp, _i, _i, _i, _i, _i64, _i64, _ovp, _vp],
    "aiko_b200_swap_rb_u8": [_vp, _vp, _i64, _vp],
    "aiko_b200_bgr2gray_u8": [_vp, _vp, _i64, _i, _vp],
    "aiko_b200_yuv_u8": [_vp, _vp, _i64, _i, _i, _vp],
    "aiko_b200_alpha_paste_u8": [_vp, _i, _i, _i, _i64, _vp, _i, _i, _i, _i, _vp],
    "aiko_b200_gray2bgr_u8": [_vp, _vp, _i64, _vp],
    "aiko_b200_flip_u8": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "aiko_b200_crop_u8": [_vp, _i, _i, _i, _i, _vp, _i, _i, _i, _i, _vp],
}

RESTYPES = {"aiko_b200_last_error": ctypes.c_char_p, "aiko_b200_resize_bicubic_workspace_bytes": ctypes.c_size_t,
            "aiko_b200_jpeg_gpu_staging_bytes": ctypes.c_size_t, "aiko_b200_jpeg_gpu_workspace_bytes": ctypes.c_size_t,
            "aiko_b200_audio_workspace_bytes": ctypes.c_size_t,
            "aiko_b200_jpeg_gpu_encode_workspace_bytes": ctypes.c_size_t,
            "aiko_b200_overlay_boxes_workspace_bytes": ctypes.c_size_t}

_lib = None


def build(verbose=False):
    """Compile libaiko_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", CSRC, "-j8"], stdout=out, stderr=None if verbose else subprocess.STDOUT)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(aiko_services_b200 has no CPU fallback)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError = header/library mismatch: fail loudly
            fn.argtypes = argtypes
            fn.restype = RESTYPES.get(name, ctypes.c_int)
        _lib = handle
    return _lib


def check(code):
    if code != OK:
        msg = lib().aiko_b200_last_error()
        raise NativeError(code, msg.decode() if msg else "")
    return code
