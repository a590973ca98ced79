"""Host -> device staging of frames: pinned buffers and cudaMemcpyAsync on a per-stream CUDA stream.

One ``FrameStager`` per aiko stream (kept in ``stream.variables``, the place the reference gives
elements for per-stream state, R/main/stream.py:81).  ``upload`` packs a list of equally shaped
frames into a pinned buffer and enqueues one asynchronous copy; buffers rotate so the host can
fill the next batch while the previous copy is in flight.
"""
import os
from contextlib import nullcontext as _nullcontext

import numpy as np
import torch

from .device_frames import DeviceBatch, DeviceImage, group_device_images


_row_runs = {}


def needed_row_runs(src_h, dst_h):
    """Runs of source rows that a bilinear resize src_h -> dst_h reads (the library's own answer,
    aiko_b200_resize_linear_rows_host), as (array('i') of first, count pairs, number of runs) -- or None
    when (nearly) every row is needed or the pattern is too fragmented to be worth strided copies."""
    key = (int(src_h), int(dst_h))
    if key not in _row_runs:
        import ctypes
        from array import array

        from . import _native
        flags = (ctypes.c_uint8 * key[0])()
        _native.check(_native.lib().aiko_b200_resize_linear_rows_host(key[0], key[1], flags))
        mask = np.frombuffer(flags, np.uint8)
        runs, y = [], 0
        while y < key[0]:
            if mask[y]:
                first = y
                while y < key[0] and mask[y]:
                    y += 1
                runs += [first, y - first]
            else:
                y += 1
        worth = mask.sum() <= 0.95 * key[0] and len(runs) // 2 <= 1024
        _row_runs[key] = (array("i", runs), len(runs) // 2) if worth else None
    return _row_runs[key]


class FrameStager:
    def __init__(self, device, slots=3, lanes=2):
        self.device = torch.device(device)
        # Consecutive batches of the stream alternate between ``lanes`` CUDA streams: frames are
        # independent, so the draw-list pass of one batch runs beside the HBM-bound resize of the next
        # (measured on the device-resident step: 45 -> 39 us with two lanes, 35 us with three).
        self.streams = [torch.cuda.Stream(self.device) for _ in range(max(lanes, 1))]
        self._lane = 0
        self.slots = slots
        self._pinned = [None] * slots
        self._events = [None] * slots
        self._next = 0
        # AIKO_B200_ZERO_COPY=1: frames written into allocate() memory stay there until their first kernel
        # and a bilinear downscale reads them over PCIe itself.  Off by default: measured on B200 the TMA
        # engine pulls 50 GB/s from host memory against 54 GB/s for the copy engine, and the 11 % of rows a
        # 1080p -> 480p resize never touches do not make up for it end to end (6.9-7.2 k vs 7.5-7.9 k frames/s).
        self.zero_copy = os.environ.get("AIKO_B200_ZERO_COPY", "0") == "1"
        self._owned = {}          # pinned blocks handed out by allocate(): base address -> tensor
        self._out = {}            # pinned result buffers by (shape): reused for device-to-host copies

    @property
    def stream(self):
        """The CUDA stream of the batch being staged (the current lane)."""
        return self.streams[self._lane]

    def next_lane(self):
        """Move on to the next CUDA stream (once per batch that enters the GPU); -> that stream."""
        self._lane = (self._lane + 1) % len(self.streams)
        return self.streams[self._lane]

    def _slot(self, nbytes):
        i = self._next
        self._next = (i + 1) % self.slots
        if self._events[i] is not None:
            self._events[i].synchronize()          # the copy out of this buffer has finished
        if self._pinned[i] is None or self._pinned[i].numel() < nbytes:
            self._pinned[i] = torch.empty(max(nbytes, 1), dtype=torch.uint8).pin_memory()
        return i, self._pinned[i]

    def allocate(self, n, frame_shape):
        """-> list of ``n`` uint8 ndarray views, each of ``frame_shape``, over one pinned buffer.

        A source element (decoder, camera reader, test generator) writes its frames here; ``upload``
        recognises such frames and lets the GPU work read them from this memory (``mapped``), with no
        intermediate host copy.  The buffer must stay untouched from the moment its frames enter the
        pipeline until their results have been read.
        """
        count = int(np.prod(frame_shape))
        pinned = torch.empty(max(n * count, 1), dtype=torch.uint8).pin_memory()
        block = pinned[:n * count].view((n,) + tuple(frame_shape))
        self._owned[block.data_ptr()] = block
        return [block[i].numpy() for i in range(n)]

    def _owned_block(self, arrays):
        """If ``arrays`` are all the frames of one ``allocate`` block, in order, return its tensor."""
        block = self._owned.get(arrays[0].ctypes.data)
        if block is None or block.shape[0] != len(arrays) or tuple(block.shape[1:]) != arrays[0].shape:
            return None
        step = block[0].numel()
        base = block.data_ptr()
        for i, a in enumerate(arrays):
            if a.ctypes.data != base + i * step or not a.flags.c_contiguous or a.dtype != np.uint8:
                return None
        return block

    def upload_array(self, frames):
        """frames: uint8 ndarray [N, H, W, C] or [N, H, W] (pageable or pinned) -> DeviceBatch"""
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        i, pinned = self._slot(frames.nbytes)
        view = pinned[:frames.nbytes].view(frames.shape)
        view.numpy()[...] = frames
        return self.upload_pinned(view, i)

    def mapped(self, pinned_tensor):
        """A caller-owned tensor in pinned host memory -> lazy DeviceBatch, not copied anywhere yet.

        If the first thing done to the frames is a bilinear downscale that reads each source row
        once, the kernel fetches the rows over PCIe itself (only the rows that carry weight, and no
        staging copy in HBM); otherwise the batch copies itself to the device when it is first used.
        Either way the work is enqueued when the frame leaves the pipeline, and the host buffer
        belongs to the caller again once the frame's results have been read.
        """
        return DeviceBatch(pinned_tensor, self.next_lane(), device=self.device)

    def upload_pinned(self, pinned_tensor, slot=None, resize_to=None):
        """A tensor already in pinned host memory -> DeviceBatch (one cudaMemcpyAsync).  ``resize_to``
        = (width, height) says that a bilinear resize is all that will read the frames: then only the
        source rows that carry weight cross PCIe (8 of 9 for 1080 -> 480; the library turns the periodic
        pattern into two strided copies per batch; AIKO_B200_COPY_NEEDED_ROWS=0 copies
        everything)."""
        lane = self.next_lane()
        runs = None
        if resize_to is not None and pinned_tensor.dim() == 4 and os.environ.get("AIKO_B200_COPY_NEEDED_ROWS", "1") == "1":
            runs = needed_row_runs(pinned_tensor.shape[1], resize_to[1])
        with torch.cuda.stream(lane):
            dev = torch.empty(pinned_tensor.shape, dtype=torch.uint8, device=self.device)
            if runs is not None:
                import ctypes

                from . import _native
                n, h, w, c = pinned_tensor.shape
                with torch.cuda.device(self.device):
                    _native.check(_native.lib().aiko_b200_copy_rows_h2d(
                        dev.data_ptr(), pinned_tensor.data_ptr(), n, h * w * c, w * c, runs[0].buffer_info()[0], runs[1],
                        ctypes.c_void_p(lane.cuda_stream)))
            else:
                dev.copy_(pinned_tensor, non_blocking=True)
            if slot is not None:
                ev = torch.cuda.Event()
                ev.record(lane)
                self._events[slot] = ev
        return DeviceBatch(dev, lane)

    def upload(self, images, resize_to=None):
        """list of ndarray / PIL.Image / DeviceImage -> list of (DeviceBatch, positions).

        Frames of equal shape are batched together; ``positions`` are their indices in ``images``.
        ``resize_to``: see ``upload_pinned`` (the caller's next step is a bilinear resize to that size).
        """
        whole = group_device_images(images)
        if whole is not None:
            return [(whole, list(range(len(images))))]
        groups = {}
        for pos, im in enumerate(images):
            if isinstance(im, DeviceImage):
                groups.setdefault(("dev", id(im.batch)), []).append(pos)
            else:
                a = im if isinstance(im, np.ndarray) else np.asarray(im)
                groups.setdefault(("host", a.shape), []).append(pos)
        out = []
        for (where, _), positions in groups.items():
            if where == "dev":
                b = images[positions[0]].batch
                idx = [images[p].index for p in positions]
                if idx == list(range(len(b))):
                    out.append((b, positions))
                else:
                    # a subset / reordering of a batch: gather ON THE BATCH'S STREAM (its frames are
                    # produced there; the gathered tensor is then owned by that stream's allocator pool)
                    source = b.tensor
                    with torch.cuda.stream(b.stream) if b.stream is not None else _nullcontext():
                        sel = torch.as_tensor(idx, device=b.device)
                        gathered = source.index_select(0, sel).contiguous()
                    sub = DeviceBatch(gathered, b.stream)
                    sub.stager = b.stager
                    out.append((sub, positions))
            else:
                arrs = [images[p] if isinstance(images[p], np.ndarray) else np.asarray(images[p]) for p in positions]
                block = self._owned_block(arrs)
                if block is not None:                 # frames were produced in our pinned memory: no host copy
                    batch = self.mapped(block) if self.zero_copy else self.upload_pinned(block, resize_to=resize_to)
                    batch.stager = self
                    out.append((batch, positions))
                    continue
                if len(arrs) > 1 and all(a.flags.c_contiguous and a.flags.writeable and a.dtype == np.uint8 for a in arrs):
                    sources = [torch.from_numpy(a) for a in arrs]     # (read-only arrays -- np.asarray(PIL image) -- are never pinned)
                    if all(t.is_pinned() for t in sources):
                        # frames of several streams, each already in page-locked memory (decoders'
                        # or cameras' own buffers): one asynchronous copy per frame straight into
                        # the batch on the device -- no staging copy on the host
                        lane = self.next_lane()
                        runs = None
                        if resize_to is not None and arrs[0].ndim == 3 and os.environ.get("AIKO_B200_COPY_NEEDED_ROWS", "1") == "1":
                            runs = needed_row_runs(arrs[0].shape[0], resize_to[1])
                        with torch.cuda.stream(lane):
                            dev = torch.empty((len(arrs),) + arrs[0].shape, dtype=torch.uint8, device=self.device)
                            if runs is not None:      # only the rows the resize reads: two copies per frame
                                import ctypes

                                from . import _native
                                h, w, c = arrs[0].shape
                                copy, st = _native.lib().aiko_b200_copy_rows_h2d, ctypes.c_void_p(lane.cuda_stream)
                                base, runs_ptr = dev.data_ptr(), runs[0].buffer_info()[0]
                                with torch.cuda.device(self.device):
                                    for k, t in enumerate(sources):
                                        _native.check(copy(base + k * h * w * c, t.data_ptr(), 1, h * w * c, w * c, runs_ptr,
                                                           runs[1], st))
                            else:
                                for k, t in enumerate(sources):
                                    dev[k].copy_(t, non_blocking=True)
                        batch = DeviceBatch(dev, lane)
                        batch.stager = self
                        out.append((batch, positions))
                        continue
                nbytes = sum(a.nbytes for a in arrs)
                i, pinned = self._slot(nbytes)
                view = pinned[:nbytes].view((len(arrs),) + arrs[0].shape)
                host = view.numpy()
                for k, a in enumerate(arrs):
                    host[k] = a
                batch = self.upload_pinned(view, i, resize_to=resize_to)
                batch.stager = self
                out.append((batch, positions))
        return out

    def pinned_out(self, shape):
        """A pinned host buffer for results of ``shape`` (reused between batches of that shape)."""
        key = tuple(shape)
        buf = self._out.get(key)
        if buf is None:
            buf = torch.empty(key, dtype=torch.uint8).pin_memory()
            self._out[key] = buf
        return buf
