"""Glyph atlas for ImageOverlay's anti-aliased labels: Python view of the atlas that lives in
libaiko_b200.so (csrc/overlay_plan.cpp, csrc/text_raster.hpp).

The reference draws labels with ``cv2.putText(image, text, (tx, ty), FONT_HERSHEY_SIMPLEX, 0.75,
(0, 255, 255), 2, cv2.LINE_AA)`` (R/elements/media/image_io.py:101-105,151-153).  The library
restates that rasteriser exactly and turns every (character, sub-pixel pen phase) into a patch
of per-pixel transfer-table ids; replaying the characters of a label in order through those
tables reproduces cv2.putText bit for bit, including labels clipped by the image border (those
patches are rasterised against the border on demand).  There is no OpenCV anywhere on this
path: the only thing taken from OpenCV is its copy of the public-domain Hershey stroke data
(csrc/hershey_simplex.inc, recovered once by tools/bake_hershey.py).

This module only mirrors the host atlas on the GPUs: the library never allocates device memory.
"""
import ctypes

import numpy as np

from . import _native
from ._native import AtlasView, check

PAD_X, PAD_Y = 8, 40          # patch origin = (floor(pen_x) - PAD_X, baseline_y - PAD_Y)
PATCH_W, PATCH_H = 40, 48
COMPONENTS = (0, 255)         # colour components of (0, 255, 255), image_io.py:101


class TextAtlas:
    """``device(dev)`` gives the GPU-resident tables; ``patch_ids`` / ``luts`` are host views."""

    _shared = None

    def __init__(self):
        self._device = {}               # torch.device -> _DeviceAtlas
        self._variants = {}             # torch.device -> ((epoch, version), variant tensors)
        self._view = AtlasView()

    @classmethod
    def shared(cls):
        if cls._shared is None:
            cls._shared = cls()
        return cls._shared

    def view(self):
        check(_native.lib().aiko_b200_overlay_atlas_host(ctypes.byref(self._view)))
        return self._view

    @property
    def version(self):
        v = self.view()
        return (v.epoch, v.version)

    @property
    def patch_ids(self):
        """uint16 [n_patches, PATCH_H, PATCH_W] over the library's memory (valid until the atlas grows)"""
        v = self.view()
        buf = (ctypes.c_uint16 * (v.n_patches * v.patch_h * v.patch_w)).from_address(v.patch_ids_host)
        return np.frombuffer(buf, np.uint16).reshape(v.n_patches, v.patch_h, v.patch_w)

    @property
    def luts(self):
        v = self.view()
        buf = (ctypes.c_uint8 * (v.n_luts * 256)).from_address(v.luts_host)
        return np.frombuffer(buf, np.uint8).reshape(v.n_luts, 256)

    def text_size(self, text):
        """cv2.getTextSize(text, FONT_HERSHEY_SIMPLEX, 0.75, 2)[0]"""
        w, h = ctypes.c_int(), ctypes.c_int()
        check(_native.lib().aiko_b200_overlay_text_size_host(_encode(text), ctypes.byref(w), ctypes.byref(h)))
        return w.value, h.value

    def variants(self, device):
        """The border variants rasterised so far as device tensors for the device-side planner
        (aiko_b200_overlay_variants_host): (keys int64 [n], firsts int32 [n], boxes int16 [n, 4], n); refreshed when
        the atlas has grown."""
        import torch
        device = torch.device(device)
        v = self.view()
        cached = self._variants.get(device)
        if cached is not None and cached[0] == (v.epoch, v.version):
            return cached[1]
        lib = _native.lib()
        n = lib.aiko_b200_overlay_variants_host(None, None, None, 0, None, None)
        keys = np.zeros(max(n, 1), np.uint64)
        firsts = np.zeros(max(n, 1), np.int32)
        boxes = np.zeros((max(n, 1), 4), np.int16)
        n = min(n, lib.aiko_b200_overlay_variants_host(keys.ctypes.data, firsts.ctypes.data, boxes.ctypes.data, len(keys), None, None))
        tensors = (torch.from_numpy(keys.view(np.int64)).to(device), torch.from_numpy(firsts).to(device),
                   torch.from_numpy(boxes).to(device), n)
        self._variants[device] = ((v.epoch, v.version), tensors)
        return tensors

    def device(self, device):
        import torch
        device = torch.device(device)
        d = self._device.get(device)
        v = self.view()
        if d is None or d.epoch != v.epoch:
            # a new epoch re-uses patch indices: kernels still queued with the old tables keep the old
            # tensors alive through their DeviceOverlay, the new epoch gets fresh ones
            d = _DeviceAtlas(v, device)
            self._device[device] = d
        d.sync(v)
        return d


def _encode(text):
    """Label bytes as cv2.putText sees them (one glyph per byte; bytes outside 32..126 draw '?')."""
    return str(text).encode("utf-8", "replace").replace(b"\0", b"?")


class _DeviceAtlas:
    """Mirror of the host atlas on one GPU: full-capacity tensors, rows appended as the host grows."""

    def __init__(self, view, device):
        import torch
        self.device = device
        self.epoch = view.epoch
        self.patch_w, self.patch_h = view.patch_w, view.patch_h
        self.patch_ids = torch.zeros((view.max_patches, view.patch_h, view.patch_w), dtype=torch.int16, device=device)
        self.luts = torch.zeros((view.max_luts, 256), dtype=torch.uint8, device=device)
        self.n_patches = self.n_luts = 0
        self.version = None
        self.ready = None            # event: the last append has been enqueued before it
        # page-locked staging for appends, allocated up front (cudaHostAlloc takes milliseconds)
        self._stage = [[torch.empty(self.STAGE_BYTES, dtype=torch.uint8).pin_memory(), None] for _ in range(4)]
        self._stage_next = 0

    STAGE_BYTES = 1 << 18

    def _push(self, dst, address, nbytes, dtype, shape, stream):
        """Host rows at ``address`` -> ``dst`` (device view).  Small appends (border variants made while
        frames are in flight) go through page-locked staging and never stall the stream; the bulk
        upload at start-up is an ordinary copy."""
        import torch
        src = np.frombuffer((ctypes.c_uint8 * nbytes).from_address(address), np.uint8)
        if nbytes > self.STAGE_BYTES:
            dst.copy_(torch.from_numpy(src.view(dtype).reshape(shape).copy()))
            return
        slot = self._stage[self._stage_next]
        self._stage_next = (self._stage_next + 1) % len(self._stage)
        if slot[1] is not None:
            slot[1].synchronize()
        slot[0][:nbytes].numpy()[...] = src
        dst.copy_(slot[0][:nbytes].view(dst.dtype).view(dst.shape), non_blocking=True)
        if slot[1] is None:
            slot[1] = torch.cuda.Event()
        slot[1].record(stream)

    def sync(self, view):
        """Copy the rows the host has appended since the last call; work that uses the tables on
        another CUDA stream waits for ``ready``."""
        import torch
        if self.version == view.version:
            return
        stream = torch.cuda.current_stream(self.device)
        if self.ready is not None:
            stream.wait_event(self.ready)
        if view.n_patches > self.n_patches:
            per = view.patch_h * view.patch_w
            rows = view.n_patches - self.n_patches
            self._push(self.patch_ids[self.n_patches:view.n_patches], view.patch_ids_host + 2 * self.n_patches * per,
                       2 * rows * per, np.int16, (rows, view.patch_h, view.patch_w), stream)
            self.n_patches = view.n_patches
        if view.n_luts > self.n_luts:
            rows = view.n_luts - self.n_luts
            self._push(self.luts[self.n_luts:view.n_luts], view.luts_host + 256 * self.n_luts, 256 * rows, np.uint8,
                       (rows, 256), stream)
            self.n_luts = view.n_luts
        self.version = view.version
        self.ready = torch.cuda.Event()
        self.ready.record(stream)
