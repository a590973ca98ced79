"""Device-resident frames handed between consecutive GPU elements.

The reference passes ``images`` -- a Python list of ndarrays / PIL images -- from element to
element through ``frame.swag`` (R/main/pipeline.py:1071).  Here a GPU element returns a list
of ``DeviceImage`` items instead: each is a view of one frame of a ``DeviceBatch`` that lives
in HBM.  A ``DeviceImage`` turns into an ndarray on demand (``np.asarray(image)``,
``image.to_pil()``), so elements that were not ported keep working, at the price of one
device-to-host copy of the batch.

A ``DeviceBatch`` is lazy: ImageResize, ImageOverlay and colour conversion only append to its
recipe.  When the pixels are needed, consecutive steps that the library can run as one kernel
(bilinear resize -> overlay -> channel swap) are launched as one.
"""
from collections import OrderedDict

import numpy as np
import torch

from . import ops
from .metrics import current_tag


_overlay_rings = OrderedDict()       # (device, CUDA stream) -> ops.OverlayRing: page-locked slots for the draw lists


class DeviceBatch:
    def __init__(self, tensor, stream=None, recipe=None, device=None, on_source_read=None, tags=None, timings=None):
        # uint8 [N, H, W, C] or [N, H, W]: on a CUDA device, or (with ``device``) in page-locked host
        # memory -- then the first kernel of the recipe reads it over PCIe if it touches every row
        # once (bilinear downscale of RGB frames), and otherwise it is copied to ``device`` first
        self._source = tensor
        self._device = torch.device(device) if device is not None else tensor.device
        self._mapped = not tensor.is_cuda
        self._on_source_read = on_source_read   # called (stream-ordered) once the host source has been consumed
        self._recipe = list(recipe or [])
        self._tags = list(tags or [None] * len(self._recipe))      # element that appended each step
        self.timings = timings if timings is not None else {}      # element -> [(start, end) events] | ("fused", element)
        self._tensor = tensor if not self._recipe and not self._mapped else None
        self._host = None
        self.stream = stream             # torch.cuda.Stream the batch's work is ordered on
        self.stager = None               # FrameStager that brought the frames in (pinned result buffers)

    # ---- geometry of the result, known without running anything
    @property
    def device(self):
        return self._device

    def __len__(self):
        return self._source.shape[0]

    @property
    def frame_shape(self):
        h, w = self._source.shape[1:3]
        c = self._source.shape[3] if self._source.dim() == 4 else 1
        for step in self._recipe:
            if step[0] in ("resize_linear", "resize_bicubic", "resize_pil"):
                w, h = step[1], step[2]
            elif step[0] == "convert":
                c = {"RGB2GRAY": 1, "BGR2GRAY": 1}.get(step[1], 3)
        return (h, w) if c == 1 else (h, w, c)

    def then(self, *step):
        tag = current_tag()
        if self._tensor is None:
            nxt = DeviceBatch(self._source, self.stream, self._recipe + [tuple(step)],
                              self._device if self._mapped else None, self._on_source_read,
                              self._tags + [tag], self.timings)
        else:
            nxt = DeviceBatch(self._tensor, self.stream, [tuple(step)], tags=[tag], timings=self.timings)
        nxt.stager = self.stager
        return nxt

    def _timed(self, first, used):
        """Start a CUDA-event pair for recipe steps [first, first + used); -> the end event to record."""
        tags = [t for t in self._tags[first:first + used] if t is not None]
        if not tags:
            return None
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        self.timings.setdefault(tags[0], []).append((start, end))
        for t in tags[1:]:
            if t != tags[0]:
                self.timings.setdefault(t, ("fused", tags[0]))
        return end

    def launch(self):
        """Enqueue the kernels of the recipe now (asynchronously); results are read later."""
        self.tensor
        return self

    # ---- materialisation
    @property
    def tensor(self):
        if self._tensor is None:
            ctx = torch.cuda.stream(self.stream) if self.stream is not None else _NullContext()
            with ctx:
                self._tensor = self._run(self._source, self._recipe)
        return self._tensor

    def _overlay(self, overlays, width, height, gray, swapped, n):
        """Plan the draw lists of this call on the host (libaiko_b200's planner: microseconds) into a page-locked
        ring slot; the bytes travel with the call that draws them (no blocking copy on the frames' stream).
        ``overlays``: one dict for all frames, or one per frame."""
        layout = "bgr" if (gray or swapped) else "rgb"
        stream = self.stream if self.stream is not None else torch.cuda.current_stream(self.device)
        key = (str(self.device), stream.cuda_stream)
        ring = _overlay_rings.get(key)
        if ring is None:
            ring = _overlay_rings[key] = ops.OverlayRing(self.device)
            if len(_overlay_rings) > 64:                  # streams come and go: forget the oldest ring
                _overlay_rings.pop(next(iter(_overlay_rings)))
        return ring.plan(overlays, width, height, layout, n_frames=n, stream=stream)

    def _source_read(self):
        if self._on_source_read is not None:
            self._on_source_read()
            self._on_source_read = None

    def _run(self, t, recipe):
        i = 0
        owned = False               # t was made by a step of this recipe: nobody else can see it, draw in place
        if self._mapped:
            first = recipe[0] if recipe else None
            direct = (first is not None and first[0] == "resize_linear" and t.dim() == 4 and
                      ops.reads_source_rows_once(tuple(t.shape[1:]), first[1], first[2]))
            if not direct:                  # everything else works on a copy in HBM
                t = t.to(self._device, non_blocking=True)
                self._source_read()
                owned = True
        while i < len(recipe):
            step = recipe[i]
            kind = step[0]
            n = t.shape[0]
            colour = t.dim() == 4 and t.shape[3] == 3
            if kind == "resize_linear":
                w, h = step[1], step[2]
                ov, swap, used = None, False, 1
                # fuse what the library runs in the same pass: [overlay] [RGB<->BGR]
                if (colour and i + used < len(recipe) and recipe[i + used][0] == "overlay"
                        and not isinstance(recipe[i + used][1], ops.DeviceDetections)):
                    ov_spec = recipe[i + used][1]
                    used += 1
                    if i + used < len(recipe) and recipe[i + used] in (("convert", "RGB2BGR"), ("convert", "BGR2RGB")):
                        swap = True
                        used += 1
                    ov = self._overlay(ov_spec, w, h, False, swap, n)
                elif colour and i + used < len(recipe) and recipe[i + used] in (("convert", "RGB2BGR"),
                                                                                 ("convert", "BGR2RGB")):
                    swap = True
                    used += 1
                end = self._timed(i, used)
                if t.is_cuda:
                    t = ops.resize_linear(t, w, h, swap_rb=swap, overlay=ov)
                else:                       # page-locked host frames, read by the kernel itself
                    t = ops.resize_linear(t, w, h, swap_rb=swap, overlay=ov, device=self._device)
                    self._source_read()
                if end is not None:
                    end.record()
                owned = True
                i += used
                continue
            end = self._timed(i, 1)
            if kind == "resize_bicubic":
                t = ops.resize_bicubic_aa(t, step[1], step[2])
            elif kind == "resize_pil":
                t = ops.resize_pil(t, step[3], step[1], step[2])
            elif kind == "overlay" and isinstance(step[1], ops.DeviceDetections):
                if not colour:
                    raise ValueError("ImageOverlay from device-resident detections needs RGB frames")
                t = ops.overlay_boxes(t, step[1], out=t if owned else torch.empty_like(t))    # planned and drawn on the GPU
            elif kind == "overlay":
                gray = t.dim() == 3
                h, w = t.shape[1:3]
                t = ops.overlay_draw(t, self._overlay(step[1], w, h, gray, False, n), out=t if owned else None)
            elif kind == "convert":
                code = step[1]
                if code in ("RGB2BGR", "BGR2RGB"):
                    t = ops.swap_rb(t)
                elif code in ("RGB2GRAY", "BGR2GRAY"):
                    t = ops.bgr2gray(t, rgb_order=(code == "RGB2GRAY"))
                elif code == "GRAY2BGR":
                    t = ops.gray2bgr(t)
                elif code in ("RGB2YUV", "BGR2YUV", "YUV2RGB", "YUV2BGR"):
                    t = ops.yuv(t, bgr_order="BGR" in code, inverse=code.startswith("YUV"))
                else:
                    raise ValueError(f"unknown conversion {code}")
            elif kind == "flip":
                t = ops.flip(t, step[1])
            elif kind == "crop":
                t = ops.crop(t, *step[1:])
            else:
                raise ValueError(f"unknown step {kind}")
            if end is not None:
                end.record()
            owned = True
            i += 1
        return t

    def host_view(self):
        """The whole batch as one ndarray over PINNED memory: one asynchronous device-to-host
        copy, no further host copy.  The stager reuses that buffer for the next result of the
        same shape on this stream, so the view is only valid until then."""
        t = self.tensor
        if self.stager is None:
            if self.stream is not None:
                self.stream.synchronize()
            return t.cpu().numpy()
        out = self.stager.pinned_out(t.shape)
        with torch.cuda.stream(self.stream):
            out.copy_(t, non_blocking=True)
        self.stream.synchronize()
        return out.numpy()

    def host(self):
        """The whole batch as one ndarray that the caller may keep (cached)."""
        if self._host is None:
            view = self.host_view()
            self._host = np.array(view, copy=True) if self.stager is not None else view
        return self._host

    def images(self):
        return [DeviceImage(self, i) for i in range(len(self))]


class _NullContext:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class DeviceImage:
    """One frame of a DeviceBatch; stands in for the ndarray the reference would pass."""

    def __init__(self, batch, index):
        self.batch, self.index = batch, index

    @property
    def shape(self):
        return self.batch.frame_shape

    @property
    def ndim(self):
        return len(self.shape)

    def __array__(self, dtype=None, copy=None):
        a = self.batch.host()[self.index]
        return a.astype(dtype) if dtype is not None and dtype != a.dtype else a

    def numpy(self):
        return np.asarray(self)

    def astype(self, dtype):
        return np.asarray(self).astype(dtype)

    pil_mode = None        # set for frames that came in as PIL images: the mode they go out with
    pil_palette = None

    def to_pil(self):
        from PIL import Image
        a = np.asarray(self)
        mode = self.pil_mode or ("L" if a.ndim == 2 else {2: "LA", 3: "RGB", 4: "RGBA"}[a.shape[2]])
        if mode == "1":
            return Image.fromarray(a > 0)
        image = Image.fromarray(a, mode)
        if mode == "P" and self.pil_palette is not None:
            image.putpalette(self.pil_palette[1], self.pil_palette[0])
        return image


def group_device_images(images):
    """If ``images`` is exactly the frames of one DeviceBatch, in order, return that batch."""
    if images and all(isinstance(im, DeviceImage) for im in images):
        b = images[0].batch
        if len(images) == len(b) and all(im.batch is b and im.index == i for i, im in enumerate(images)):
            return b
    return None
