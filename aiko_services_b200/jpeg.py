"""JPEG at the file edges of the path, on the GPU: the decode half of ``PIL.Image.open(path)`` ...
``image.resize(...)`` (R/elements/media/image_io.py:184,218) and the encode inside ``image.save(path)``
(image_io.py:250), so that frames are born and leave device-resident instead of crossing PCIe as raw pixels.

``GpuJpegDecoder`` / ``GpuJpegEncoder``: this library's own kernels (csrc/jpeg_decode.cu, csrc/jpeg_encode.cu in
libaiko_b200.so; include/aiko_b200.h): baseline files, pixels and file bytes identical to Pillow's.
``JpegDecoder``: nvJPEG behind libaiko_b200_jpeg.so (include/aiko_b200_jpeg.h) for what the own decoder refuses
(progressive files).  torch only provides the tensors and the streams; there is no CPU decoder here."""
import ctypes
import os

import torch

from . import _native
from ._native import NativeError

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaiko_b200_jpeg.so")
_vp, _i, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
SIGNATURES = {
    "aiko_b200_jpeg_decoder_create": [ctypes.POINTER(_vp), _i, _i],
    "aiko_b200_jpeg_decoder_destroy": [_vp],
    "aiko_b200_jpeg_decoder_is_hardware": [_vp],
    "aiko_b200_jpeg_last_error": [],
    "aiko_b200_jpeg_info": [_vp, ctypes.c_char_p, _sz, ctypes.POINTER(_i), ctypes.POINTER(_i), ctypes.POINTER(_i),
                            ctypes.POINTER(_i)],
    "aiko_b200_jpeg_decode_rgb_u8": [_vp, _vp, _vp, _i, _vp, _i, _i, ctypes.c_int64, _vp],
}
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C aiko_services_b200/csrc` "
                              "(there is no CPU decoder behind this module)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.argtypes = argtypes
            fn.restype = ctypes.c_char_p if name.endswith("last_error") else ctypes.c_int
        _lib = handle
    return _lib


def _check(code):
    if code != 0:
        msg = lib().aiko_b200_jpeg_last_error()
        raise NativeError(code, msg.decode() if msg else "")


def is_jpeg(data):
    return len(data) > 3 and data[0] == 0xFF and data[1] == 0xD8 and data[2] == 0xFF


class JpegDecoder:
    """Batched JPEG -> uint8 [n, h, w, 3] RGB CUDA tensor.  One decoder per (GPU, thread)."""

    def __init__(self, device, max_batch=64, backend=0):
        self.device = torch.device(device)
        self.max_batch = max_batch
        self._handle = _vp()
        with torch.cuda.device(self.device):
            _check(lib().aiko_b200_jpeg_decoder_create(ctypes.byref(self._handle), backend, max_batch))
        self.hardware = bool(lib().aiko_b200_jpeg_decoder_is_hardware(self._handle))

    def info(self, data):
        """-> (width, height, components, subsampling) from the JPEG header"""
        w, h, c, s = _i(), _i(), _i(), _i()
        _check(lib().aiko_b200_jpeg_info(self._handle, data, len(data), ctypes.byref(w), ctypes.byref(h),
                                         ctypes.byref(c), ctypes.byref(s)))
        return w.value, h.value, c.value, s.value

    def decode(self, buffers, out=None, stream=None):
        """``buffers``: list of bytes-like JPEG images of one size (bytes, or uint8 tensors in pinned
        memory).  -> uint8 tensor [n, h, w, 3] on the device; work is enqueued on ``stream``."""
        n = len(buffers)
        if n == 0:
            return torch.empty((0, 0, 0, 3), dtype=torch.uint8, device=self.device)
        ptrs, lens, keep = (_vp * n)(), (_sz * n)(), []
        for k, b in enumerate(buffers):
            if isinstance(b, torch.Tensor):
                ptrs[k], lens[k] = b.data_ptr(), b.numel()
            else:
                raw = bytes(b) if not isinstance(b, bytes) else b
                keep.append(raw)
                ptrs[k] = ctypes.cast(ctypes.c_char_p(raw), _vp).value
                lens[k] = len(raw)
        first = buffers[0]
        head = bytes(first[:min(len(first), 1 << 16)].numpy().tobytes()) if isinstance(first, torch.Tensor) else keep[0]
        w, h, _, _ = self.info(head)
        if out is None:
            out = torch.empty((n, h, w, 3), dtype=torch.uint8, device=self.device)
        elif not (out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and tuple(out.shape) == (n, h, w, 3)):
            raise RuntimeError(f"jpeg decode out: a contiguous uint8 CUDA tensor of shape {(n, h, w, 3)} is needed, "
                               f"got {tuple(out.shape)}")
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            for start in range(0, n, self.max_batch):
                count = min(self.max_batch, n - start)
                _check(lib().aiko_b200_jpeg_decode_rgb_u8(
                    self._handle, ctypes.byref(ptrs, start * ctypes.sizeof(_vp)), ctypes.byref(lens, start * ctypes.sizeof(_sz)),
                    count, out[start].data_ptr(), h, w, h * w * 3, _vp(s.cuda_stream)))
        return out

    def close(self):
        if self._handle:
            lib().aiko_b200_jpeg_decoder_destroy(self._handle)
            self._handle = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_decoders = {}


def decoder_for(device, max_batch=64):
    key = (str(torch.device(device)), max_batch)
    if key not in _decoders:
        _decoders[key] = JpegDecoder(device, max_batch)
    return _decoders[key]


class GpuJpegDecoder:
    """This library's own JPEG decoder (csrc/jpeg_decode.cu; include/aiko_b200.h, aiko_b200_jpeg_gpu_*):
    Huffman decoding, IDCT, upsampling and colour conversion all on the GPU, pixels bit-identical to
    Pillow's decoder.  Baseline files (gray, 4:4:4, 4:2:2, 4:2:0), with or without restart markers; for
    anything else (progressive, arithmetic, 12 bit) ``decode`` raises NativeError(ERR_UNSUPPORTED) and the
    caller keeps its other decoder (``JpegDecoder`` above, or Pillow).  One decoder per (GPU, CUDA stream)."""

    def __init__(self, device):
        self.device = torch.device(device)
        self._buffers = {}        # geometry key -> (workspace, staging, capacity, segments)
        self._event = None        # the staging buffer's last upload has run
        self._last = None

    @staticmethod
    def info(data):
        """-> (width, height, components, hmax, vmax, restart_interval) from the JPEG headers"""
        vals = [_i() for _ in range(6)]
        raw = data if isinstance(data, bytes) else bytes(data)
        _native.check(_native.lib().aiko_b200_jpeg_gpu_info_host(raw, len(raw), *[ctypes.byref(v) for v in vals]))
        return tuple(v.value for v in vals)

    @classmethod
    def _header(cls, buffer):
        """``info`` of a file given as bytes or as a uint8 host tensor (only as much of a tensor is copied as the
        header needs: 64 KB, or more when EXIF / ICC segments push the frame header further back)."""
        if not isinstance(buffer, torch.Tensor):
            return cls.info(bytes(buffer))
        size = 1 << 16
        while True:
            size = min(size, buffer.numel())
            try:
                return cls.info(buffer[:size].numpy().tobytes())
            except NativeError as error:
                if size == buffer.numel() or error.code == _native.ERR_UNSUPPORTED:
                    raise
                size *= 16

    def _pointers(self, buffers):
        n = len(buffers)
        ptrs, lens, keep = (_vp * n)(), (_sz * n)(), []
        for k, b in enumerate(buffers):
            if isinstance(b, torch.Tensor):
                ptrs[k], lens[k] = b.data_ptr(), b.numel()
            else:
                raw = b if isinstance(b, bytes) else bytes(b)
                keep.append(raw)
                ptrs[k] = ctypes.cast(ctypes.c_char_p(raw), _vp).value
                lens[k] = len(raw)
        return ptrs, lens, keep

    def decode(self, buffers, out=None, stream=None, check=False):
        """``buffers``: JPEG images of one geometry (bytes, or uint8 tensors in page-locked memory).
        -> uint8 tensor [n, h, w, 3] RGB on the device, enqueued on ``stream``.  ``check`` waits for the
        result and raises if the entropy-coded data was corrupt."""
        n = len(buffers)
        if n == 0:
            return torch.empty((0, 0, 0, 3), dtype=torch.uint8, device=self.device)
        w, h, comps, hmax, vmax, restart = self._header(buffers[0])
        mcus = -(-w // (8 * hmax)) * -(-h // (8 * vmax))
        segments = -(-mcus // restart) if restart > 0 else 1
        ptrs, lens, keep = self._pointers(buffers)
        total = sum(int(v) for v in lens)
        key = (n, w, h, comps, hmax, vmax)
        lib = _native.lib()
        if out is None:
            out = torch.empty((n, h, w, 3), dtype=torch.uint8, device=self.device)
        elif not (out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and tuple(out.shape) == (n, h, w, 3)):
            raise RuntimeError(f"jpeg decode out: a contiguous uint8 CUDA tensor of shape {(n, h, w, 3)} is needed, "
                               f"got {tuple(out.shape)}")
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        if self._event is not None:
            self._event.synchronize()            # the previous call has left the staging buffer and the workspace
        for attempt in range(2):
            entry = self._buffers.get(key)
            if entry is None or entry[2] < total or entry[3] < segments:
                capacity = max(-(-total * 5 // 4 // (1 << 20)) * (1 << 20), 1 << 20)
                nbytes = lib.aiko_b200_jpeg_gpu_workspace_bytes(n, w, h, comps, hmax, vmax, capacity, segments)
                if nbytes == 0:
                    raise NativeError(_native.ERR_UNSUPPORTED, f"jpeg: unsupported layout ({comps} components, {hmax}x{vmax})")
                with torch.cuda.device(self.device):
                    workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
                staging = torch.empty(lib.aiko_b200_jpeg_gpu_staging_bytes(n), dtype=torch.uint8).pin_memory()
                entry = self._buffers[key] = (workspace, staging, capacity, segments)
            workspace, staging, capacity, max_segments = entry
            with torch.cuda.device(self.device):
                rc = lib.aiko_b200_jpeg_gpu_decode_rgb_u8(
                    ptrs, lens, n, out.data_ptr(), h, w, h * w * 3, staging.data_ptr(), staging.numel(), workspace.data_ptr(),
                    workspace.numel(), capacity, max_segments, _vp(s.cuda_stream))
            if rc != _native.ERR_WORKSPACE or attempt:
                _native.check(rc)
                break
            # the other images of the batch have more restart segments than the first: size for the largest
            for k in range(n):
                restart_k = self._header(buffers[k])[5]
                segments = max(segments, -(-mcus // restart_k) if restart_k > 0 else 1)
        if self._event is None:
            self._event = torch.cuda.Event()
        self._event.record(s)
        self._last = (workspace, n, w, h, comps, hmax, vmax, capacity, max_segments, s, keep)
        if check:
            self.check()
        return out

    def check(self):
        """Wait for the last decode and raise NativeError if its entropy-coded data was corrupt."""
        if self._last is not None:
            workspace, n, w, h, comps, hmax, vmax, capacity, segs, s, _ = self._last
            with torch.cuda.device(self.device):
                _native.check(_native.lib().aiko_b200_jpeg_gpu_status(workspace.data_ptr(), n, w, h, comps, hmax, vmax,
                                                                      capacity, segs, _vp(s.cuda_stream)))


class GpuJpegEncoder:
    """This library's JPEG encoder (csrc/jpeg_encode.cu; include/aiko_b200.h, aiko_b200_jpeg_gpu_encode_*):
    colour conversion, downsampling, forward DCT, quantisation and Huffman coding on the GPU; the files are
    byte-identical to what ``PIL.Image.fromarray(frame).save(f, "JPEG", quality=..., subsampling=...)`` writes.
    One encoder per (GPU, CUDA stream)."""

    SUBSAMPLING = {0: (1, 1), 1: (2, 1), 2: (2, 2), "4:4:4": (1, 1), "4:2:2": (2, 1), "4:2:0": (2, 2)}

    def __init__(self, device):
        self.device = torch.device(device)
        self._buffers = {}        # geometry key -> (workspace, out, sizes, scan_capacity, out_stride)
        self._room = {}           # geometry key -> scan capacity as a multiple of half the raw frame

    def _key(self, frames, subsampling):
        gray = frames.dim() == 3
        hs, vs = subsampling if isinstance(subsampling, tuple) else self.SUBSAMPLING[subsampling]
        return (frames.shape[0], frames.shape[1], frames.shape[2], 1 if gray else frames.shape[3], hs, vs)

    def encode_async(self, frames, quality=75, subsampling=2, stream=None):
        """``frames``: uint8 CUDA tensor [n, h, w, 3] (RGB) or [n, h, w] (gray).  Enqueues the encoder on
        ``stream``; -> (out uint8 [n, out_stride] on the device: file k = out[k, :sizes[k]]; sizes int32 [n] on
        the device, -1 = did not fit).  Both belong to the encoder and are overwritten by its next call."""
        if not (isinstance(frames, torch.Tensor) and frames.is_cuda and frames.dtype == torch.uint8 and frames.is_contiguous()):
            raise RuntimeError("GpuJpegEncoder: frames must be a contiguous uint8 CUDA tensor")
        key = self._key(frames, subsampling)
        n, h, w, channels, hs, vs = key
        if channels not in (1, 3):
            raise NativeError(_native.ERR_UNSUPPORTED, f"jpeg encode: {channels} channels")
        lib = _native.lib()
        entry = self._buffers.get(key)
        if entry is None:
            scan_capacity = max(h * w * channels // 2 * self._room.get(key, 1), 4096)
            out_stride = -(-(scan_capacity + scan_capacity // 16 + 1024) // 16) * 16
            nbytes = lib.aiko_b200_jpeg_gpu_encode_workspace_bytes(n, w, h, channels, hs, vs, scan_capacity)
            if nbytes == 0:
                raise NativeError(_native.ERR_UNSUPPORTED, f"jpeg encode: unsupported layout ({channels} channels, {hs}x{vs})")
            with torch.cuda.device(self.device):
                workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
                out = torch.empty((n, out_stride), dtype=torch.uint8, device=self.device)
                sizes = torch.empty(n, dtype=torch.int32, device=self.device)
            entry = self._buffers[key] = (workspace, out, sizes, scan_capacity, out_stride)
        workspace, out, sizes, scan_capacity, out_stride = entry
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        from .torch_ops import OPS
        OPS.jpeg_encode_u8(frames, int(quality), hs, vs, workspace, scan_capacity, out, sizes, s.cuda_stream)
        return out, sizes

    def encode(self, frames, quality=75, subsampling=2, stream=None):
        """-> list of ``bytes``, one JPEG file per frame (waits for the result)."""
        if frames.shape[0] == 0:
            return []
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        key = self._key(frames, subsampling)
        while True:
            out, sizes = self.encode_async(frames, quality, subsampling, stream=s)
            with torch.cuda.stream(s):
                host_sizes = sizes.cpu()
            s.synchronize()
            lengths = host_sizes.tolist()
            if min(lengths) >= 0:
                break
            # a scan did not fit (noise at high quality): four times the room, up to four raw frames
            room = self._room.get(key, 1) * 4
            if room > 16:
                raise NativeError(_native.ERR_WORKSPACE, "jpeg encode: the compressed data does not fit its buffers")
            self._room[key] = room
            del self._buffers[key]
        longest = max(lengths)
        with torch.cuda.stream(s):
            host = out[:, :longest].cpu()
        s.synchronize()
        data = host.numpy()
        return [data[k, :lengths[k]].tobytes() for k in range(len(lengths))]


_gpu_encoders = {}


def gpu_encoder_for(device):
    key = str(torch.device(device))
    if key not in _gpu_encoders:
        _gpu_encoders[key] = GpuJpegEncoder(device)
    return _gpu_encoders[key]


_gpu_decoders = {}


def gpu_decoder_for(device):
    key = str(torch.device(device))
    if key not in _gpu_decoders:
        _gpu_decoders[key] = GpuJpegDecoder(device)
    return _gpu_decoders[key]


def decode_on_gpu(buffers, device, stream=None, check=False):
    """JPEG images of one size -> uint8 [n, h, w, 3] RGB on ``device``: this library's own decoder when the
    files allow it (baseline; bit-identical to Pillow), nvJPEG otherwise.  ``check`` waits for the result and
    raises on corrupt entropy-coded data (own decoder).  -> (tensor, "aiko" | "nvjpeg")"""
    try:
        return gpu_decoder_for(device).decode(buffers, stream=stream, check=check), "aiko"
    except NativeError as error:
        if error.code != _native.ERR_UNSUPPORTED:
            raise
    return decoder_for(device).decode(buffers, stream=stream), "nvjpeg"
