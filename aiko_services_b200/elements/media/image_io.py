"""GPU image elements with the reference's names, inputs, outputs and parameters
(R/elements/media/image_io.py).  ``process_frame`` only enqueues work: frames are staged to the
GPU once (pinned memory + cudaMemcpyAsync on the stream's CUDA stream), stay device-resident
between consecutive elements as lazy ``DeviceImage`` items, and consecutive resize / overlay /
colour-convert steps run as one fused kernel when the pixels are finally needed.

``logic_*`` functions hold the behaviour; the classes below bind them to this package's
runtime, and ``aiko_services_b200.reference_adapter`` binds the same functions to the
reference's own ``aiko.PipelineElement`` so unmodified pipelines pick them up.
"""
import numpy as np

from ...device_frames import DeviceImage
from ...pipeline import PipelineElement
from ...staging import FrameStager
from ...stream import StreamEvent
from .common_io import DataSource, DataTarget, contains_all

__all__ = ["ImageConvert", "ImageOutput", "ImageOverlay", "ImageReadFile", "ImageResize", "ImageToTensor",
           "ImageWriteFile"]

_CONVERSIONS = ("RGB2BGR", "BGR2RGB", "RGB2GRAY", "BGR2GRAY", "GRAY2BGR", "RGB2YUV", "BGR2YUV", "YUV2RGB", "YUV2BGR")
_FLIPS = {"vertical": 0, "horizontal": 1, "both": -1}


def device_for(stream_id, n_devices=None):
    """Streams are pinned to one GPU for their lifetime: stable hash of stream_id."""
    import torch
    import zlib
    n = n_devices or torch.cuda.device_count()
    if n <= 0:
        raise RuntimeError("aiko_services_b200 needs a CUDA device (there is no CPU fallback)")
    current = torch.cuda.current_device()
    if n_devices is None and "LOCAL_RANK" in __import__("os").environ:
        return current          # one process per GPU: the launcher already chose
    return zlib.crc32(str(stream_id).encode()) % n


def stager_for(stream):
    """Per-stream CUDA stream + pinned staging ring, kept where the reference keeps per-stream
    state (stream.variables, R/main/stream.py:81)."""
    stager = stream.variables.get("_b200_stager")
    if stager is None:
        import torch
        stager = FrameStager(torch.device("cuda", device_for(stream.stream_id)))
        stream.variables["_b200_stager"] = stager
    return stager


def _apply(stream, images, step):
    """Append one recipe step to every batch behind ``images`` -> new list, same order."""
    if not images:
        return []
    out = [None] * len(images)
    hint = (step[1], step[2]) if step[0] == "resize_linear" else None     # only the rows a resize reads are staged
    for batch, positions in stager_for(stream).upload(images, resize_to=hint):
        result = batch.then(*step).images()
        for k, pos in enumerate(positions):
            out[pos] = result[k]
    return out


# ---- behaviour shared by both bindings ------------------------------------------------------
def logic_image_resize(element, stream, images):
    """ImageResize.process_frame (image_io.py:208-220): ndarray frames take cv2's bilinear
    arithmetic, PIL frames take what PIL.Image.resize does for their mode (antialiased bicubic;
    RGBA / LA premultiplied around it; P / 1 nearest)."""
    from ... import ops
    resolution, _ = element.get_parameter("resolution")
    width, height = resolution.split("x")
    width, height = int(width), int(height)
    out = [None] * len(images)
    arrays = [i for i, im in enumerate(images) if isinstance(im, (np.ndarray, DeviceImage)) and not isinstance(im, _LazyPil)]
    if arrays:
        for i, r in zip(arrays, _apply(stream, [images[i] for i in arrays], ("resize_linear", width, height))):
            out[i] = r
    by_mode = {}
    for i, im in enumerate(images):
        if out[i] is None and i not in arrays:
            mode = getattr(im, "pil_mode", None) or getattr(im, "mode", None)
            if mode not in ops.PIL_PLAIN_MODES + ops.PIL_ALPHA_MODES + ops.PIL_NEAREST_MODES:
                return StreamEvent.ERROR, {"diagnostic": f"ImageResize: PIL mode {mode!r} is not supported on the GPU "
                                                         "(8-bit modes L, LA, RGB, RGBA, CMYK, YCbCr, HSV, LAB, P, 1)"}
            by_mode.setdefault(mode, []).append(i)
    for mode, idx in by_mode.items():
        frames = []
        for i in idx:
            im = images[i]
            if isinstance(im, _LazyPil):
                frames.append(im)
            else:
                a = np.asarray(im)
                frames.append(a.astype(np.uint8) * 255 if mode == "1" else a)
        for i, r in zip(idx, _apply(stream, frames, ("resize_pil", width, height, mode))):
            palette = getattr(images[i], "pil_palette", None)
            if palette is None and mode == "P":
                palette = (images[i].palette.mode, images[i].getpalette(None))
            out[i] = _LazyPil(r, mode, palette)      # the reference returns PIL images for PIL inputs
    return StreamEvent.OKAY, {"images": out}


class _LazyPil(DeviceImage):
    """A DeviceImage that came from a PIL input: ``save`` / ``size`` / ``mode`` like the PIL image the
    reference would have produced, pixels still on the GPU until asked for."""

    def __init__(self, image, mode="RGB", palette=None):
        super().__init__(image.batch, image.index)
        self.pil_mode, self.pil_palette = mode, palette

    @property
    def mode(self):
        return self.pil_mode

    @property
    def size(self):
        return (self.shape[1], self.shape[0])

    def save(self, path, *a, **k):
        return self.to_pil().save(path, *a, **k)


def logic_image_overlay(element, stream, images, overlay):
    """ImageOverlay.process_frame (image_io.py:107-163).  ``overlay``: the reference's dict, or -- from a detector
    element that keeps its output on the GPU -- an ``ops.DeviceDetections``, planned and drawn on the device."""
    for image in images:
        if not isinstance(image, (np.ndarray, DeviceImage)) or isinstance(image, _LazyPil):
            # the reference indexes image.shape, which PIL images lack -> its frame errors out
            return StreamEvent.ERROR, {"diagnostic": "ImageOverlay: images must be numpy arrays"}
    return StreamEvent.OKAY, {"images": _apply(stream, images, ("overlay", overlay))}


def logic_image_convert(element, stream, images):
    conversion, found = element.get_parameter("conversion", default=None)
    flip, flip_found = element.get_parameter("flip", default=None)
    if found:
        if conversion not in _CONVERSIONS:
            return StreamEvent.ERROR, {"diagnostic": f"ImageConvert: unknown conversion {conversion}"}
        images = _apply(stream, images, ("convert", conversion))
    if flip_found:
        if flip not in _FLIPS:
            return StreamEvent.ERROR, {"diagnostic": f"ImageConvert: unknown flip {flip}"}
        images = _apply(stream, images, ("flip", _FLIPS[flip]))
    return StreamEvent.OKAY, {"images": images}


def logic_image_to_tensor(element, stream, images):
    """BASELINE config 4: frames -> the input a torch model element expects, without leaving the GPU:
    cv2-exact bilinear resize to ``resolution`` and ((x / 255) - mean) / std as float32 [N, 3, H, W].
    Parameters: ``resolution`` "WxH", ``mean`` and ``std`` "r,g,b" (default: the ImageNet constants).
    Output ``tensor``: a CUDA tensor a downstream element can pass straight to its model (the reference's
    detectors hand ndarrays to ultralytics / deepface instead, R/examples/yolo/yolo.py:62-64)."""
    from ... import ops
    resolution, _ = element.get_parameter("resolution")
    width, height = (int(v) for v in resolution.split("x"))
    mean, _ = element.get_parameter("mean", default="0.485,0.456,0.406")
    std, _ = element.get_parameter("std", default="0.229,0.224,0.225")
    mean = [float(v) for v in str(mean).split(",")]
    std = [float(v) for v in str(std).split(",")]
    if not images:
        return StreamEvent.OKAY, {"tensor": None}
    parts = [None] * len(images)
    for batch, positions in stager_for(stream).upload(images):
        frames = batch.tensor
        if frames.dim() != 4 or frames.shape[3] != 3:
            return StreamEvent.ERROR, {"diagnostic": "ImageToTensor: RGB frames expected"}
        import torch
        with torch.cuda.stream(batch.stream) if batch.stream is not None else _null():
            out = ops.resize_norm_f32(frames, width, height, mean, std)
        for k, pos in enumerate(positions):
            parts[pos] = out[k]
    import torch
    return StreamEvent.OKAY, {"tensor": torch.stack(parts) if len(parts) > 1 else parts[0][None]}


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def logic_image_read_file(element, stream, paths):
    """ImageReadFile.process_frame (image_io.py:180-191): lazy PIL open, decode deferred into the
    consumer, exactly as the reference -- unless the parameter ``decode`` is "gpu": then baseline
    colour baseline JPEG files are decoded on the GPU (csrc/jpeg_decode.cu through aiko_services_b200.jpeg) and
    handed on as device-resident frames that downstream elements treat as the RGB PIL images the
    reference would have produced (ImageResize takes its PIL branch).  Other files stay lazy PIL."""
    from PIL import Image
    decode, _ = element.get_parameter("decode", default="host")
    images = [None] * len(paths)
    if str(decode).lower() == "gpu":
        try:
            _read_jpegs_on_gpu(stream, paths, images)
        except Exception as exception:
            return StreamEvent.ERROR, {"diagnostic": f"Error decoding image: {exception}"}
    for k, path in enumerate(paths):
        if images[k] is None:
            try:
                images[k] = Image.open(path)
            except Exception as exception:
                return StreamEvent.ERROR, {"diagnostic": f"Error loading image: {exception}"}
    return StreamEvent.OKAY, {"images": images}


def _read_jpegs_on_gpu(stream, paths, images):
    """Fills ``images[k]`` with a device-resident frame for every path that is a colour JPEG;
    equal-sized files are decoded as one batch."""
    from ... import jpeg
    from ...device_frames import DeviceBatch
    stager = stager_for(stream)
    groups = {}
    for k, path in enumerate(paths):
        with open(path, "rb") as f:
            data = f.read()
        if not jpeg.is_jpeg(data):
            continue
        try:
            width, height, components, hmax, vmax, restart = jpeg.GpuJpegDecoder.info(data)
        except Exception:
            continue                      # progressive, CMYK, 12-bit, ...: the reference's host path
        groups.setdefault((width, height, components, hmax, vmax), []).append((k, data))
    import torch
    for key, members in groups.items():
        lane = stager.next_lane()
        with torch.cuda.stream(lane):
            # this library's decoder (bit-identical to Pillow); a truncated or corrupt file is an error of this
            # element, as a failing Image.open / load is in the reference, not a frame of garbage
            frames, _ = jpeg.decode_on_gpu([data for _, data in members], stager.device, stream=lane, check=True)
            mode = "RGB"
            if key[2] == 1:               # a gray file opens as mode L in Pillow: one channel (the decoder wrote R = G = B)
                frames, mode = frames[..., 0].contiguous(), "L"
        batch = DeviceBatch(frames, lane)
        batch.stager = stager
        for image, (k, _) in zip(batch.images(), members):
            images[k] = _LazyPil(image, mode)


_JPEG_SUFFIXES = (".jpg", ".jpeg", ".jpe", ".jfif")


def logic_image_write_file(element, stream, images):
    """ImageWriteFile.process_frame (image_io.py:234-255).  With the parameter ``encode: gpu``, device-resident
    RGB / gray frames that go to a JPEG path are encoded on the GPU (aiko_services_b200.jpeg.GpuJpegEncoder:
    the bytes image.save(path) would write) and only the compressed files cross PCIe."""
    from PIL import Image
    encode, _ = element.get_parameter("encode", default="host")
    paths = []
    for image in images:
        path = stream.variables["target_path"]
        if contains_all(path, "{}"):
            path = path.format(stream.variables["target_file_id"])
            stream.variables["target_file_id"] += 1
        paths.append(path)
    written = set()
    if str(encode).lower() == "gpu":
        try:
            written = _write_jpegs_on_gpu(images, paths)
        except Exception as exception:
            return StreamEvent.ERROR, {"diagnostic": f"Error saving image: {exception}"}
    for k, (image, path) in enumerate(zip(images, paths)):
        if k in written:
            continue
        if isinstance(image, DeviceImage):
            image = image.to_pil()
        elif not isinstance(image, Image.Image):
            if isinstance(image, np.ndarray):
                image = Image.fromarray(image.astype("uint8"), "RGB")
            else:
                return StreamEvent.ERROR, {"diagnostic": "UNKNOWN IMAGE TYPE"}
        try:
            image.save(path)
        except Exception as exception:
            return StreamEvent.ERROR, {"diagnostic": f"Error saving image: {exception}"}
    return StreamEvent.OKAY, {}


def _write_jpegs_on_gpu(images, paths):
    """Encodes the device-resident RGB / L frames whose path is a JPEG path, one batch per DeviceBatch, and
    writes the files.  -> indices written."""
    import torch
    from ... import jpeg
    groups = {}
    for k, (image, path) in enumerate(zip(images, paths)):
        if not isinstance(image, DeviceImage) or not str(path).lower().endswith(_JPEG_SUFFIXES):
            continue
        shape = image.shape
        mode = image.pil_mode or ("L" if len(shape) == 2 else "RGB" if shape[2] == 3 else None)
        if mode not in ("RGB", "L"):
            continue
        groups.setdefault(id(image.batch), (image.batch, []))[1].append((k, image.index))
    written = set()
    for batch, members in groups.values():
        frames = batch.tensor
        indices = [i for _, i in members]
        lane = batch.stream if batch.stream is not None else torch.cuda.current_stream(frames.device)
        with torch.cuda.stream(lane):
            if indices != list(range(frames.shape[0])):
                frames = frames.index_select(0, torch.as_tensor(indices, device=frames.device))
            files = jpeg.gpu_encoder_for(frames.device).encode(frames.contiguous(), stream=lane)
        for (k, _), data in zip(members, files):
            with open(paths[k], "wb") as f:
                f.write(data)
            written.add(k)
    return written


def _pipeline_stager(pipeline, device_index):
    """Stager shared by the streams of one GPU, used when frames of many streams travel together."""
    import torch
    stagers = pipeline.__dict__.setdefault("_b200_stagers", {})
    if device_index not in stagers:
        stagers[device_index] = FrameStager(torch.device("cuda", device_index), slots=3)
    return stagers[device_index]


def _cross_stream(prepare):
    """Build ``process_frame_batch`` for an element whose work is one recipe step per frame.

    ``prepare(element, stream, **inputs)`` returns either ``(StreamEvent.ERROR, diagnostic)`` or
    ``(images, step, merge_key)``.  Frames of all calls that share a GPU and a merge key (and a
    shape) are staged together and extended by ONE step, so the kernel that eventually runs
    covers every concurrent stream -- batching keyed by stream_id as the frames' owners.
    ``step`` may end in a per-call argument (ImageOverlay's dict); it is expanded per frame.
    """
    def process_frame_batch(self, calls, pipeline):
        results = [None] * len(calls)
        buckets = {}                       # (device, merge_key) -> [(call index, images, step)]
        for ci, (stream, inputs) in enumerate(calls):
            pipeline._enable_thread_local(stream, stream.frame_id)
            try:
                prepared = prepare(self, stream, **inputs)
            except Exception as exception:      # same containment as a per-frame call
                prepared = (StreamEvent.ERROR, {"diagnostic": repr(exception)})
            finally:
                pipeline._disable_thread_local()
            if len(prepared) == 2:
                results[ci] = prepared
                continue
            images, step, merge_key = prepared
            buckets.setdefault((device_for(stream.stream_id), merge_key), []).append((ci, images, step))
        for (device_index, _), members in buckets.items():
            flat, owner = [], []
            for ci, images, step in members:
                for k, image in enumerate(images):
                    flat.append(image)
                    owner.append((ci, k, step))
            outputs = {ci: [None] * len(images) for ci, images, _ in members}
            if flat:
                first_step = owner[0][2]
                hint = (first_step[1], first_step[2]) if first_step[0] == "resize_linear" else None
                for batch, positions in _pipeline_stager(pipeline, device_index).upload(flat, resize_to=hint):
                    step = owner[positions[0]][2]
                    if step[0] == "overlay":   # one overlay dict per frame, in batch order
                        step = ("overlay", [owner[p][2][1] for p in positions])
                    for image, pos in zip(batch.then(*step).images(), positions):
                        ci, k, _ = owner[pos]
                        outputs[ci][k] = image
            for ci, _, _ in members:
                results[ci] = (StreamEvent.OKAY, {"images": outputs[ci]})
        return results
    return process_frame_batch


def _prepare_resize(element, stream, images):
    if any(not isinstance(im, (np.ndarray, DeviceImage)) or isinstance(im, _LazyPil) for im in images):
        return logic_image_resize(element, stream, images)       # PIL frames: per-stream path
    resolution, _ = element.get_parameter("resolution")
    width, height = (int(v) for v in resolution.split("x"))
    return images, ("resize_linear", width, height), ("resize_linear", width, height)


def _prepare_overlay(element, stream, images, overlay):
    for image in images:
        if not isinstance(image, (np.ndarray, DeviceImage)) or isinstance(image, _LazyPil):
            return StreamEvent.ERROR, {"diagnostic": "ImageOverlay: images must be numpy arrays"}
    from ... import ops
    if isinstance(overlay, ops.DeviceDetections):
        return logic_image_overlay(element, stream, images, overlay)    # device-resident detections: per-stream path
    return images, ("overlay", overlay), "overlay"


def _prepare_convert(element, stream, images):
    conversion, found = element.get_parameter("conversion", default=None)
    flip, flip_found = element.get_parameter("flip", default=None)
    if flip_found or not found or conversion not in _CONVERSIONS:
        return logic_image_convert(element, stream, images)      # rare combinations: per-stream path
    return images, ("convert", conversion), ("convert", conversion)


# ---- this package's runtime bindings -------------------------------------------------------------
class ImageOutput(PipelineElement):
    def process_frame(self, stream, images):
        return StreamEvent.OKAY, {"images": images}


class ImageResize(PipelineElement):
    batch_across_streams = True
    process_frame = logic_image_resize
    process_frame_batch = _cross_stream(_prepare_resize)


class ImageOverlay(PipelineElement):
    batch_across_streams = True
    process_frame = logic_image_overlay
    process_frame_batch = _cross_stream(_prepare_overlay)


class ImageConvert(PipelineElement):
    """GPU home of the colour-conversion / flip call sites of R/elements/media/video_io.py:129,299
    and webcam_io.py:122-129 (those elements' capture and codec parts are out of scope).
    Parameters: ``conversion`` in RGB2BGR | BGR2RGB | RGB2GRAY | BGR2GRAY | GRAY2BGR and / or
    ``flip`` in horizontal | vertical | both."""
    batch_across_streams = True
    process_frame = logic_image_convert
    process_frame_batch = _cross_stream(_prepare_convert)


class ImageToTensor(PipelineElement):
    process_frame = logic_image_to_tensor


class ImageReadFile(DataSource):
    process_frame = logic_image_read_file


class ImageWriteFile(DataTarget):
    process_frame = logic_image_write_file
