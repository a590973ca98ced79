"""The C ABI of libaiko_b200.so as PyTorch custom ops: ``torch.ops.aiko_b200.*``.

BASELINE.json's north_star asks for "a thin C-ABI layer exposed as PyTorch custom ops"; this module
is that layer.  Every op is one entry point of include/aiko_b200.h: tensors carry the device
pointers and shapes, outputs are passed in (``Tensor(a!) out`` -- nothing allocates), ``stream`` is a
cudaStream_t handle as an integer (``torch.cuda.current_stream().cuda_stream``; 0 = the legacy
default stream) and ``overlay`` is the address of an ``aiko_b200_overlay`` struct in host memory
(``ops.DeviceOverlay.struct``; 0 = none).  Errors of the library surface as ``NativeError``.
``aiko_services_b200.ops`` / ``audio_ops`` are written on top of these ops.

There is no CPU implementation: the ops only call the library.
"""
import ctypes

import torch

from . import _native
from ._native import check

_LIB = torch.library.Library("aiko_b200", "DEF")
_vp = ctypes.c_void_p


def _op(schema):
    def register(fn):
        name = schema.split("(", 1)[0]
        _LIB.define(schema)
        _LIB.impl(name, fn, "CompositeExplicitAutograd")
        return fn
    return register


def _geometry(t):
    """uint8 frames [N, H, W, C] or [N, H, W] -> (n, h, w, c)"""
    if t.dtype != torch.uint8 or not t.is_contiguous() or not (t.is_cuda or t.is_pinned()):
        raise ValueError("aiko_b200 ops take contiguous uint8 frames on a CUDA device (or in page-locked host memory)")
    if t.dim() == 4:
        return t.shape[0], t.shape[1], t.shape[2], t.shape[3]
    if t.dim() == 3:
        return t.shape[0], t.shape[1], t.shape[2], 1
    raise ValueError(f"expected frames [N,H,W,C] or [N,H,W], got {tuple(t.shape)}")


def _current(t):
    """Run the call with t's device current (cheap when it already is)."""
    from .ops import _on_device
    return _on_device(t.device)


@_op("resize_linear_u8(Tensor src, Tensor(a!) out, bool swap_rb, int overlay, int stream) -> ()")
def resize_linear_u8(src, out, swap_rb, overlay, stream):
    n, h, w, c = _geometry(src)
    _, oh, ow, _ = _geometry(out)
    with _current(out):
        check(_native.lib().aiko_b200_resize_linear_u8(
            src.data_ptr(), n, h, w, c, w * c, h * w * c, out.data_ptr(), oh, ow, ow * c, oh * ow * c,
            int(swap_rb), ctypes.cast(overlay, _native._ovp) if overlay else None, _vp(stream)))


@_op("resize_bicubic_prepare(Tensor(a!) workspace, int src_h, int src_w, int dst_h, int dst_w, int stream) -> ()")
def resize_bicubic_prepare(workspace, src_h, src_w, dst_h, dst_w, stream):
    with _current(workspace):
        check(_native.lib().aiko_b200_resize_bicubic_prepare(workspace.data_ptr(), workspace.numel(), src_h, src_w,
                                                             dst_h, dst_w, _vp(stream)))


@_op("resize_bicubic_aa_u8(Tensor src, Tensor(a!) out, Tensor workspace, int stream) -> ()")
def resize_bicubic_aa_u8(src, out, workspace, stream):
    n, h, w, c = _geometry(src)
    _, oh, ow, _ = _geometry(out)
    with _current(src):
        check(_native.lib().aiko_b200_resize_bicubic_aa_u8(src.data_ptr(), n, h, w, c, out.data_ptr(), oh, ow,
                                                           workspace.data_ptr(), workspace.numel(), _vp(stream)))


@_op("resize_bicubic_aa_u8_noalloc(Tensor src, Tensor(a!) out, int stream) -> ()")
def resize_bicubic_aa_u8_noalloc(src, out, stream):
    n, h, w, c = _geometry(src)
    _, oh, ow, _ = _geometry(out)
    with _current(src):
        check(_native.lib().aiko_b200_resize_bicubic_aa_u8_noalloc(src.data_ptr(), n, h, w, c, out.data_ptr(), oh, ow,
                                                                   _vp(stream)))


@_op("resize_norm_f32_nchw(Tensor src, Tensor(a!) out, float[] mean, float[] std, int stream) -> ()")
def resize_norm_f32_nchw(src, out, mean, std, stream):
    n, h, w, _ = _geometry(src)
    m = (ctypes.c_float * 3)(*[float(v) for v in mean])
    s = (ctypes.c_float * 3)(*[float(v) for v in std])
    with _current(src):
        check(_native.lib().aiko_b200_resize_norm_f32_nchw(src.data_ptr(), n, h, w, out.data_ptr(), out.shape[2],
                                                           out.shape[3], m, s, _vp(stream)))


@_op("overlay_u8(Tensor src, Tensor(a!) out, int overlay, int stream) -> ()")
def overlay_u8(src, out, overlay, stream):
    n, h, w, c = _geometry(src)
    with _current(src):
        check(_native.lib().aiko_b200_overlay_u8(src.data_ptr(), out.data_ptr(), n, h, w, c, w * c, h * w * c,
                                                 ctypes.cast(overlay, _native._ovp), _vp(stream)))


@_op("overlay_boxes_u8(Tensor(a!) frames, int layout, Tensor boxes, Tensor class_ids, Tensor confidences, Tensor names, "
     "int n_names, Tensor patch_ids, Tensor luts, int n_patches, Tensor variant_keys, Tensor variant_firsts, "
     "Tensor variant_boxes, int n_variants, Tensor workspace, Tensor(b!) status, int stream) -> ()")
def overlay_boxes_u8(frames, layout, boxes, class_ids, confidences, names, n_names, patch_ids, luts, n_patches, variant_keys,
                     variant_firsts, variant_boxes, n_variants, workspace, status, stream):
    n, h, w, _ = _geometry(frames)
    with _current(frames):
        check(_native.lib().aiko_b200_overlay_boxes_u8(
            frames.data_ptr(), n, h, w, h * w * 3, int(layout), boxes.data_ptr(), class_ids.data_ptr(), confidences.data_ptr(),
            boxes.shape[0], boxes.shape[1], names.data_ptr(), int(n_names), names.shape[1], patch_ids.data_ptr(),
            luts.data_ptr(), int(n_patches), variant_keys.data_ptr(), variant_firsts.data_ptr(), variant_boxes.data_ptr(),
            int(n_variants), workspace.data_ptr(), workspace.numel(), status.data_ptr(), _vp(stream)))


@_op("swap_rb_u8(Tensor src, Tensor(a!) out, int stream) -> ()")
def swap_rb_u8(src, out, stream):
    with _current(src):
        check(_native.lib().aiko_b200_swap_rb_u8(src.data_ptr(), out.data_ptr(), src.numel() // 3, _vp(stream)))


@_op("bgr2gray_u8(Tensor src, Tensor(a!) out, bool rgb_order, int stream) -> ()")
def bgr2gray_u8(src, out, rgb_order, stream):
    with _current(src):
        check(_native.lib().aiko_b200_bgr2gray_u8(src.data_ptr(), out.data_ptr(), src.numel() // 3, int(rgb_order),
                                                  _vp(stream)))


@_op("yuv_u8(Tensor src, Tensor(a!) out, bool bgr_order, bool inverse, int stream) -> ()")
def yuv_u8(src, out, bgr_order, inverse, stream):
    with _current(src):
        check(_native.lib().aiko_b200_yuv_u8(src.data_ptr(), out.data_ptr(), src.numel() // 3, int(bgr_order),
                                             int(inverse), _vp(stream)))


@_op("alpha_paste_u8(Tensor(a!) frames, Tensor rgba, int x, int y, int stream) -> ()")
def alpha_paste_u8(frames, rgba, x, y, stream):
    n, h, w, _ = _geometry(frames)
    with _current(frames):
        check(_native.lib().aiko_b200_alpha_paste_u8(frames.data_ptr(), n, h, w, h * w * 3, rgba.data_ptr(), rgba.shape[0],
                                                     rgba.shape[1], int(x), int(y), _vp(stream)))


@_op("gray2bgr_u8(Tensor src, Tensor(a!) out, int stream) -> ()")
def gray2bgr_u8(src, out, stream):
    with _current(src):
        check(_native.lib().aiko_b200_gray2bgr_u8(src.data_ptr(), out.data_ptr(), src.numel(), _vp(stream)))


@_op("flip_u8(Tensor src, Tensor(a!) out, int flip_code, int stream) -> ()")
def flip_u8(src, out, flip_code, stream):
    n, h, w, c = _geometry(src)
    with _current(src):
        check(_native.lib().aiko_b200_flip_u8(src.data_ptr(), out.data_ptr(), n, h, w, c, int(flip_code), _vp(stream)))


@_op("crop_u8(Tensor src, Tensor(a!) out, int x0, int y0, int stream) -> ()")
def crop_u8(src, out, x0, y0, stream):
    n, h, w, c = _geometry(src)
    _, oh, ow, _ = _geometry(out)
    with _current(src):
        check(_native.lib().aiko_b200_crop_u8(src.data_ptr(), n, h, w, c, out.data_ptr(), int(x0), int(y0), ow, oh,
                                              _vp(stream)))


@_op("premultiply_u8(Tensor src, Tensor(a!) out, int channels, bool inverse, int stream) -> ()")
def premultiply_u8(src, out, channels, inverse, stream):
    with _current(src):
        check(_native.lib().aiko_b200_premultiply_u8(src.data_ptr(), out.data_ptr(), src.numel() // src.shape[-1],
                                                     int(channels), int(inverse), _vp(stream)))


@_op("gather_nearest_u8(Tensor src, Tensor(a!) out, Tensor xtab, Tensor ytab, int stream) -> ()")
def gather_nearest_u8(src, out, xtab, ytab, stream):
    n, h, w, c = _geometry(src)
    _, oh, ow, _ = _geometry(out)
    with _current(src):
        check(_native.lib().aiko_b200_gather_nearest_u8(src.data_ptr(), n, h, w, c, out.data_ptr(), oh, ow,
                                                        xtab.data_ptr(), ytab.data_ptr(), _vp(stream)))


@_op("audio_prepare(Tensor(a!) workspace, int n_samples, int stream) -> ()")
def audio_prepare(workspace, n_samples, stream):
    with _current(workspace):
        check(_native.lib().aiko_b200_audio_prepare(workspace.data_ptr(), workspace.numel(), n_samples, _vp(stream)))


@_op("audio_chain(Tensor audio, float sample_rate, Tensor(a!) amplitudes, Tensor(b!)? band_sums, int[] band_ranges, "
     "Tensor(c!)? top_freq, Tensor(d!)? top_amp, Tensor(e!)? top_count, int gate_lo, int gate_hi, Tensor workspace, "
     "int stream) -> ()")
def audio_chain(audio, sample_rate, amplitudes, band_sums, band_ranges, top_freq, top_amp, top_count, gate_lo, gate_hi,
                workspace, stream):
    ranges = (ctypes.c_int * 16)(*band_ranges) if len(band_ranges) == 16 else None
    ptr = lambda t: t.data_ptr() if t is not None else None
    with _current(audio):
        check(_native.lib().aiko_b200_audio_chain(
            audio.data_ptr(), 0 if audio.dtype == torch.int16 else 1, audio.shape[0], audio.shape[1], float(sample_rate),
            amplitudes.data_ptr(), ptr(band_sums), ranges, ptr(top_freq), ptr(top_amp), ptr(top_count), int(gate_lo),
            int(gate_hi), workspace.data_ptr(), workspace.numel(), _vp(stream)))


@_op("audio_filter_f64(Tensor amplitudes, Tensor frequencies, Tensor(a!) top_freq, Tensor(b!) top_amp, "
     "Tensor(c!) top_count, int stream) -> ()")
def audio_filter_f64(amplitudes, frequencies, top_freq, top_amp, top_count, stream):
    b, n = amplitudes.shape
    with _current(amplitudes):
        check(_native.lib().aiko_b200_audio_filter_f64(amplitudes.data_ptr(), frequencies.data_ptr(), b, n,
                                                       0 if frequencies.shape[0] == 1 else n, top_freq.data_ptr(),
                                                       top_amp.data_ptr(), top_count.data_ptr(), _vp(stream)))


@_op("audio_bands_f64(Tensor amplitudes, Tensor frequencies, Tensor(a!) band_freq, Tensor(b!) band_amp, int stream) -> ()")
def audio_bands_f64(amplitudes, frequencies, band_freq, band_amp, stream):
    b, n = amplitudes.shape
    with _current(amplitudes):
        check(_native.lib().aiko_b200_audio_bands_f64(amplitudes.data_ptr(), frequencies.data_ptr(), b, n,
                                                      0 if frequencies.shape[0] == 1 else n, band_freq.data_ptr(),
                                                      band_amp.data_ptr(), _vp(stream)))


@_op("jpeg_encode_u8(Tensor src, int quality, int hs, int vs, Tensor workspace, int scan_capacity, Tensor(a!) out, "
     "Tensor(b!) sizes, int stream) -> ()")
def jpeg_encode_u8(src, quality, hs, vs, workspace, scan_capacity, out, sizes, stream):
    n, h, w, c = _geometry(src)
    with _current(src):
        check(_native.lib().aiko_b200_jpeg_gpu_encode_u8(
            src.data_ptr(), n, h, w, c, h * w * c, int(quality), int(hs), int(vs), workspace.data_ptr(), workspace.numel(),
            int(scan_capacity), out.data_ptr(), out.shape[1], sizes.data_ptr(), _vp(stream)))


OPS = torch.ops.aiko_b200
