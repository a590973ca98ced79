"""GPU: the reference-facing element / pipeline layer.  Existing pipeline-definition JSON
(module names of the reference) runs on this package's runtime, frames stay device-resident
between elements, frames of many streams are batched into one launch, and -- where the
reference package is present -- the same GPU elements run inside the reference's own,
unmodified PipelineImpl.  Results are compared with the reference elements' cv2 / Pillow calls
(oracle/reference_port.py) and the C oracle."""
import json
import os
import queue

import numpy as np
import pytest
import torch

from aiko_services_b200.device_frames import DeviceImage
from aiko_services_b200.pipeline import Pipeline
from golden import cases
from oracle import image as O
from oracle import ref_harness, reference_port

pytestmark = pytest.mark.gpu

IMAGE_IO = "aiko_services.elements.media.image_io"     # what existing definitions name

SOURCE = '''
from aiko_services_b200.pipeline import PipelineElement
from aiko_services_b200.stream import StreamEvent

class Feed(PipelineElement):
    def process_frame(self, stream, **inputs):
        return StreamEvent.OKAY, inputs
'''


def definition(tmp_path, names, resolution="64x48", extra=None):
    src = tmp_path / "feed.py"
    src.write_text(SOURCE)
    io = {"Feed": (["images", "overlay"], ["images", "overlay"]), "ImageResize": (["images"], ["images"]),
          "ImageOverlay": (["images", "overlay"], ["images"]), "ImageConvert": (["images"], ["images"]),
          "ImageOutput": (["images"], ["images"])}
    elements = []
    for name in names:
        i, o = io[name]
        e = {"name": name, "input": [{"name": k, "type": "any"} for k in i],
             "output": [{"name": k, "type": "any"} for k in o],
             "deploy": {"local": {"module": str(src) if name == "Feed" else IMAGE_IO}}}
        if extra and name in extra:
            e["parameters"] = extra[name]
        elements.append(e)
    return {"version": 0, "name": "p_gpu", "runtime": "python", "graph": ["(" + " ".join(names) + ")"],
            "parameters": {"resolution": resolution}, "elements": elements}


def test_chain_device_resident_and_fused(cuda, tmp_path):
    names = ["Feed", "ImageResize", "ImageOverlay", "ImageConvert", "ImageOutput"]
    p = Pipeline(definition(tmp_path, names, "160x120", {"ImageConvert": {"conversion": "RGB2BGR"}}))
    frames = [cases.frame(700 + k, 270, 480, 3) for k in range(4)]
    q = queue.Queue()
    assert p.create_stream("s1", queue_response=q)
    okay, swag = p.process_frame({"stream_id": "s1", "frame_id": 0}, {"images": frames, "overlay": cases.OVERLAY_SMALL})
    assert okay
    images = swag["images"]
    assert all(isinstance(im, DeviceImage) for im in images)
    batch = images[0].batch
    # one lazy recipe for the three elements (run as one C-ABI call), enqueued when the frame left the pipeline
    assert [s[0] for s in batch._recipe] == ["resize_linear", "overlay", "convert"] and batch._tensor is not None
    ref = reference_port.chain_resize_overlay_bgr(frames, "160x120", cases.OVERLAY_SMALL)
    for im, r in zip(images, ref):
        np.testing.assert_array_equal(np.asarray(im), r)          # D2H on demand, like an ndarray
    assert images[0].shape == (120, 160, 3)


def test_frames_in_pinned_memory_are_read_by_the_kernel_itself(cuda, tmp_path, monkeypatch):
    """Frames written into FrameStager.allocate() memory: a downscale reads them over PCIe with no
    staging copy, anything else copies them first; both equal the reference arithmetic."""
    from aiko_services_b200.elements.media.image_io import stager_for
    names = ["Feed", "ImageResize", "ImageOverlay", "ImageConvert", "ImageOutput"]
    src = [cases.frame(900 + k, 270, 480, 3) for k in range(3)]
    for zero_copy, resolution in (("1", "160x120"), ("0", "160x120"), ("1", "400x250"), ("1", "640x360")):
        monkeypatch.setenv("AIKO_B200_ZERO_COPY", zero_copy)
        p = Pipeline(definition(tmp_path, names, resolution, {"ImageConvert": {"conversion": "RGB2BGR"}}))
        assert p.create_stream("s1")
        views = stager_for(p.streams["s1"]).allocate(3, (270, 480, 3))
        for v, f in zip(views, src):
            v[...] = f
        okay, swag = p.process_frame({"stream_id": "s1", "frame_id": 0}, {"images": views, "overlay": cases.OVERLAY_SMALL})
        assert okay
        batch = swag["images"][0].batch
        w, h = (int(v) for v in resolution.split("x"))
        direct = zero_copy == "1" and 270 >= 2 * h             # each needed source row is fetched once
        assert batch._mapped == (zero_copy == "1") and batch._source.is_cuda == (zero_copy == "0")
        assert direct == (zero_copy == "1" and ops_reads_once((270, 480, 3), w, h))
        ref = reference_port.chain_resize_overlay_bgr(src, resolution, cases.OVERLAY_SMALL)
        for im, r in zip(swag["images"], ref):
            np.testing.assert_array_equal(np.asarray(im), r)
    # an overlay-only pipeline (no resize first) copies the pinned frames, then draws
    p = Pipeline(definition(tmp_path, ["Feed", "ImageOverlay", "ImageOutput"], "480x270"))
    assert p.create_stream("s2")
    views = stager_for(p.streams["s2"]).allocate(3, (270, 480, 3))
    for v, f in zip(views, src):
        v[...] = f
    okay, swag = p.process_frame({"stream_id": "s2", "frame_id": 0}, {"images": views, "overlay": cases.OVERLAY_SMALL})
    assert okay
    for im, r in zip(swag["images"], reference_port.image_overlay(src, cases.OVERLAY_SMALL)):
        np.testing.assert_array_equal(np.asarray(im), r)


def ops_reads_once(shape, w, h):
    from aiko_services_b200 import ops
    return ops.reads_source_rows_once(shape, w, h)


def test_cross_stream_batching_one_launch(cuda, tmp_path):
    p = Pipeline(definition(tmp_path, ["Feed", "ImageResize", "ImageOverlay", "ImageOutput"], "96x64"))
    rng = np.random.default_rng(1)
    work, expect = [], []
    for k in range(12):
        sid = f"s{k:02d}"
        assert p.create_stream(sid)
        frame = cases.frame(800 + k, 128, 192, 3)
        overlay = {"objects": [{"name": f"id{k}", "confidence": 0.5 + k / 100}],
                   "rectangles": [{"x": 5 + k, "y": 30, "w": 40, "h": 20}]}
        work.append(({"stream_id": sid, "frame_id": 0}, {"images": [frame], "overlay": overlay}))
        expect.append(reference_port.image_overlay(reference_port.image_resize([frame], "96x64"), overlay)[0])
    results = p.process_frames(work)
    batches = {id(swag["images"][0].batch): str(swag["images"][0].batch.device) for okay, swag in results}
    # 12 streams -> one DeviceBatch -> one fused launch per GPU the streams are pinned to (one on a 1-GPU box)
    assert len(batches) == len(set(batches.values())) <= torch.cuda.device_count()
    for (okay, swag), ref in zip(results, expect):
        assert okay
        np.testing.assert_array_equal(np.asarray(swag["images"][0]), ref)


def test_pil_branch_read_resize_write(cuda, tmp_path):
    from PIL import Image
    for k in range(3):
        Image.fromarray(cases.frame(900 + k, 120, 160, 3), "RGB").save(tmp_path / f"in_{k:02d}.png")
    d = {"version": 0, "name": "p_image_0", "runtime": "python",
         "graph": ["(ImageReadFile ImageResize ImageWriteFile)"], "parameters": {"resolution": "80x60"},
         "elements": [
             {"name": "ImageReadFile", "parameters": {"data_sources": f"(file://{tmp_path}/in_{{}}.png)",
                                                      "data_batch_size": 2},
              "input": [{"name": "paths", "type": "[Path]"}], "output": [{"name": "images", "type": "[image]"}],
              "deploy": {"local": {"module": IMAGE_IO}}},
             {"name": "ImageResize", "input": [{"name": "images", "type": "[image]"}],
              "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": IMAGE_IO}}},
             {"name": "ImageWriteFile", "parameters": {"data_targets": f"file://{tmp_path}/out_{{:02d}}.png"},
              "input": [{"name": "images", "type": "[image]"}], "output": [],
              "deploy": {"local": {"module": IMAGE_IO}}}]}
    p = Pipeline(d)
    assert p.create_stream("1")
    p.run()
    for k in range(3):
        got = np.asarray(Image.open(tmp_path / f"out_{k:02d}.png"))
        ref = np.asarray(Image.open(tmp_path / f"in_{k:02d}.png").resize((80, 60)))
        np.testing.assert_array_equal(got, ref)


def test_errors_become_stream_events(cuda, tmp_path):
    from PIL import Image
    p = Pipeline(definition(tmp_path, ["Feed", "ImageOverlay", "ImageOutput"]))
    assert p.create_stream("e")
    pil = Image.fromarray(cases.frame(1, 32, 32, 3), "RGB")
    okay, _ = p.process_frame({"stream_id": "e"}, {"images": [pil], "overlay": {}})
    assert not okay and "e" not in p.streams          # ERROR destroys the stream, as in the reference
    p = Pipeline(definition(tmp_path, ["Feed", "ImageResize", "ImageOutput"], resolution="bogus"))
    assert p.create_stream("f")
    okay, _ = p.process_frame({"stream_id": "f"}, {"images": [cases.frame(1, 32, 32, 3)], "overlay": {}})
    assert not okay and "f" not in p.streams


def test_gray_and_mixed_shapes(cuda, tmp_path):
    p = Pipeline(definition(tmp_path, ["Feed", "ImageResize", "ImageOverlay", "ImageOutput"], "64x48"))
    assert p.create_stream("g")
    frames = [cases.frame(1, 96, 128, 3), cases.frame(2, 96, 128, 1), cases.frame(3, 50, 70, 3)]
    okay, swag = p.process_frame({"stream_id": "g"}, {"images": frames, "overlay": cases.OVERLAY_RECTS_ONLY})
    assert okay
    ref = reference_port.image_overlay(reference_port.image_resize(frames, "64x48"), cases.OVERLAY_RECTS_ONLY)
    for im, r in zip(swag["images"], ref):
        np.testing.assert_array_equal(np.asarray(im), r)


def test_image_to_tensor_feeds_a_torch_model_element(cuda, tmp_path):
    """BASELINE config 4 shape: frames -> fp32 NCHW CUDA tensor -> a torch module, no host round trip."""
    model_src = tmp_path / "model.py"
    model_src.write_text('''
import torch
from aiko_services_b200.pipeline import PipelineElement
from aiko_services_b200.stream import StreamEvent

class TinyModel(PipelineElement):
    def process_frame(self, stream, tensor):
        assert tensor.is_cuda and tensor.dtype == torch.float32
        return StreamEvent.OKAY, {"scores": tensor.mean(dim=(2, 3))}
''')
    d = {"version": 0, "name": "p_cfg4", "runtime": "python", "graph": ["(ImageToTensor TinyModel)"],
         "parameters": {"resolution": "56x56"},
         "elements": [{"name": "ImageToTensor", "input": [{"name": "images", "type": "[image]"}],
                       "output": [{"name": "tensor", "type": "tensor"}], "deploy": {"local": {"module": IMAGE_IO}}},
                      {"name": "TinyModel", "input": [{"name": "tensor", "type": "tensor"}],
                       "output": [{"name": "scores", "type": "tensor"}], "deploy": {"local": {"module": str(model_src)}}}]}
    p = Pipeline(d)
    assert p.create_stream("m")
    frames = [cases.frame(950 + k, 270, 480, 3) for k in range(3)]
    okay, swag = p.process_frame({"stream_id": "m"}, {"images": frames})
    assert okay and tuple(swag["tensor"].shape) == (3, 3, 56, 56)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    for k in range(3):
        ref = O.resize_norm_f32(frames[k], 56, 56, mean, std)
        np.testing.assert_array_equal(swag["tensor"][k].cpu().numpy(), ref)
        np.testing.assert_allclose(swag["scores"][k].cpu().numpy(), ref.mean(axis=(1, 2)), rtol=1e-5)


@pytest.mark.skipif(not ref_harness.available(), reason="reference package not present (baseline/install_ref.sh)")
def test_drop_in_inside_the_unmodified_reference_runtime(cuda):
    """The reference's own PipelineImpl, its own JSON module names, GPU elements underneath."""
    ref_harness.load_reference()
    from aiko_services_b200 import reference_adapter
    module = reference_adapter.install()
    try:
        frames = [cases.frame(2000 + i, 1080, 1920, 3) for i in range(2)]
        out = ref_harness.run_reference_pipeline(
            [("ImageResize", IMAGE_IO, ["images"], ["images"], {"resolution": "640x480"}),
             ("ImageOverlay", IMAGE_IO, ["images", "overlay"], ["images"], None),
             ("ImageOutput", IMAGE_IO, ["images"], ["images"], None)],
            [{"images": frames, "overlay": cases.OVERLAY_4}])
        images = out[0]["images"]
        assert type(images[0]).__name__ == "DeviceImage"          # the GPU classes were the ones loaded
        ref = reference_port.image_overlay(reference_port.image_resize(frames, "640x480"), cases.OVERLAY_4)
        for im, r in zip(images, ref):
            np.testing.assert_array_equal(np.asarray(im), r)
    finally:
        reference_adapter.uninstall()


@pytest.mark.gpu
def test_subset_of_a_lazy_batch_is_gathered_on_its_own_stream(cuda):
    """An element that forwards only some images of a lazy batch: the gather must run after (and on
    the stream of) the kernels that produce the frames, for many rounds without a synchronize."""
    import numpy as np
    from aiko_services_b200.staging import FrameStager
    from golden import cases
    from oracle import image as O
    stager = FrameStager(cuda)
    frames = [cases.frame(3100 + k, 540, 960, 3) for k in range(6)]
    want = [O.resize_linear(f, 320, 180)[:, :, ::-1] for f in frames]
    for _ in range(20):
        (batch, _), = stager.upload(frames)
        resized = batch.then("resize_linear", 320, 180).images()
        picked = [resized[4], resized[1], resized[5]]
        (sub, positions), = stager.upload(picked)
        out = sub.then("convert", "RGB2BGR").host()
        for k, src in zip(positions, (4, 1, 5)):
            np.testing.assert_array_equal(out[k], want[src])


@pytest.mark.gpu
def test_pil_rgba_and_palette_images_through_the_elements(cuda, tmp_path):
    """ImageReadFile -> ImageResize -> ImageWriteFile on RGBA and palette PNGs: the files written equal
    what Pillow's own resize gives (mode kept, alpha premultiplied around the filter, P nearest)."""
    from PIL import Image
    for mode in ("RGBA", "P", "LA"):
        src = cases.pil_image(4242, mode, 120, 200)
        src.save(tmp_path / f"in_{mode}_0.png")
        d = {"version": 0, "name": "p_png", "runtime": "python", "graph": ["(ImageReadFile ImageResize ImageWriteFile)"],
             "parameters": {"resolution": "64x48"},
             "elements": [
                 {"name": "ImageReadFile", "input": [{"name": "paths", "type": "[path]"}],
                  "output": [{"name": "images", "type": "[image]"}],
                  "parameters": {"data_sources": f"(file://{tmp_path}/in_{mode}_{{}}.png)", "data_batch_size": 1},
                  "deploy": {"local": {"module": IMAGE_IO}}},
                 {"name": "ImageResize", "input": [{"name": "images", "type": "[image]"}],
                  "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": IMAGE_IO}}},
                 {"name": "ImageWriteFile", "input": [{"name": "images", "type": "[image]"}], "output": [],
                  "parameters": {"data_targets": f"file://{tmp_path}/out_{mode}.png"},
                  "deploy": {"local": {"module": IMAGE_IO}}}]}
        p = Pipeline(d)
        assert p.create_stream("1")
        p.run()
        got = Image.open(tmp_path / f"out_{mode}.png")
        want = Image.open(tmp_path / f"in_{mode}_0.png").resize((64, 48))
        assert got.mode == want.mode == mode
        np.testing.assert_array_equal(np.asarray(got.convert("RGBA")), np.asarray(want.convert("RGBA")))


@pytest.mark.gpu
def test_element_metrics_carry_gpu_time(cuda, tmp_path, monkeypatch):
    """frame.metrics["pipeline_elements"] (R/main/pipeline.py:1170-1182; consumer PE_Metrics,
    R/examples/pipeline/elements.py:133-149): GPU elements report host time + CUDA-event time of the
    kernels they caused; an element fused into another's kernel says so."""
    import aiko_services_b200.pipeline as P
    from aiko_services_b200.metrics import GpuSeconds
    captured = {}

    class KeepFrame(P.Frame):          # the runtime releases a Frame when it completes: keep a handle
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            captured["frame"] = self
    monkeypatch.setattr(P, "Frame", KeepFrame)
    p = Pipeline(definition(tmp_path, ["Feed", "ImageResize", "ImageOverlay", "ImageOutput"], "320x180"))
    assert p.create_stream("m")
    frames = [cases.frame(5200 + k, 540, 960, 3) for k in range(4)]
    for frame_id in range(2):
        okay, swag = p.process_frame({"stream_id": "m", "frame_id": frame_id},
                                     {"images": frames, "overlay": cases.OVERLAY_SMALL})
        assert okay
    elements = captured["frame"].metrics["pipeline_elements"]
    resize, overlay = elements["time_ImageResize"], elements["time_ImageOverlay"]
    assert isinstance(resize, GpuSeconds) and isinstance(overlay, GpuSeconds)
    assert overlay.fused_into == "ImageResize" and not overlay.events
    assert 2e-6 < resize.gpu < 0.05                                    # microseconds .. milliseconds of kernels
    assert float(resize) >= resize.gpu and f"{resize * 1000:.3f}"      # what PE_Metrics does with the value
    assert not isinstance(elements["time_Feed"], GpuSeconds)


@pytest.mark.gpu
def test_staging_only_the_rows_a_resize_reads(cuda, monkeypatch):
    """AIKO_B200_COPY_NEEDED_ROWS=1: frames staged for a bilinear resize cross PCIe without the source rows
    that carry no weight (aiko_b200_resize_linear_rows_host / aiko_b200_copy_rows_h2d); same pixels."""
    from aiko_services_b200.staging import FrameStager, needed_row_runs
    runs = needed_row_runs(1080, 480)
    assert runs is not None and sum(list(runs[0])[1::2]) == 960 and needed_row_runs(480, 480) is None
    monkeypatch.setenv("AIKO_B200_COPY_NEEDED_ROWS", "1")
    stager = FrameStager(cuda)
    frames = [cases.frame(6100 + k, 270, 480, 3) for k in range(3)]
    (batch, _), = stager.upload(frames, resize_to=(160, 120))
    out = batch.then("resize_linear", 160, 120).host()
    for k in range(3):
        np.testing.assert_array_equal(out[k], O.resize_linear(frames[k], 160, 120))


def test_needed_rows_staging_random_shapes(cuda):
    """The stager ships only the rows a bilinear resize reads, as strided copies when the pattern is periodic and
    as one whole copy when it is too fragmented (aiko_b200_copy_rows_h2d): random source / destination heights,
    batches, both entry routes (one pinned block; one pinned buffer per frame).  The device buffer is poisoned
    first, so a row that should have been copied and was not would show."""
    from aiko_services_b200.staging import FrameStager
    rng = np.random.default_rng(77)
    stager = FrameStager(cuda)
    for trial in range(24):
        sh, dh = int(rng.integers(8, 400)), int(rng.integers(1, 200))
        if trial % 4 == 0:
            dh = max(1, sh * 4 // 9)                       # the periodic 2.25x pattern of the headline shape
        sw, dw, n = int(rng.integers(4, 200)), int(rng.integers(1, 100)), int(rng.integers(1, 5))
        frames = [cases.frame(7000 + 10 * trial + k, sh, sw, 3) for k in range(n)]
        poison = torch.full((n, sh, sw, 3), 0xA5, dtype=torch.uint8, device=cuda)      # next torch.empty of this size
        del poison
        if trial % 2:
            views = stager.allocate(n, (sh, sw, 3))        # one pinned block owned by the stager
            for v, f in zip(views, frames):
                v[...] = f
            (batch, _), = stager.upload(views, resize_to=(dw, dh))
        else:
            pinned = []
            for f in frames:                               # one pinned buffer per frame (frames of several streams)
                t = torch.empty(f.shape, dtype=torch.uint8).pin_memory()
                t.numpy()[...] = f
                pinned.append(t.numpy())
            (batch, _), = stager.upload(pinned, resize_to=(dw, dh))
        out = batch.then("resize_linear", dw, dh).host()
        for k in range(n):
            np.testing.assert_array_equal(out[k], O.resize_linear(frames[k], dw, dh), err_msg=f"trial {trial}: {sw}x{sh} -> {dw}x{dh}")


def test_image_overlay_element_takes_device_resident_detections(cuda, tmp_path):
    """A detector element that keeps its output on the GPU hands ImageOverlay an ops.DeviceDetections instead of
    the overlay dict: ImageResize -> ImageOverlay through the pipeline, same pixels as with the dict."""
    from aiko_services_b200 import ops
    names = ["person", "car"]
    boxes = torch.tensor([[[30, 40, 60, 30], [90, 60, 40, 30], [0, 0, 0, 0]]], dtype=torch.int32, device=cuda)
    cls = torch.tensor([[0, 1, 0]], dtype=torch.int32, device=cuda)
    conf = torch.tensor([[0.875, 0.33, 0.0]], dtype=torch.float32, device=cuda)
    det = ops.DeviceDetections(boxes, cls, conf, names)
    frames = [cases.frame(9500 + k, 270, 480, 3) for k in range(2)]
    results = []
    for overlay in (det, det.to_overlay_dicts()[0]):
        p = Pipeline(definition(tmp_path, ["Feed", "ImageResize", "ImageOverlay", "ImageOutput"], "240x135"))
        assert p.create_stream("s")
        okay, swag = p.process_frame({"stream_id": "s", "frame_id": 0}, {"images": list(frames), "overlay": overlay})
        assert okay
        results.append([np.asarray(im) for im in swag["images"]])
    for a, b in zip(*results):
        np.testing.assert_array_equal(a, b)
    ref = reference_port.image_overlay(reference_port.image_resize(frames, "240x135"), det.to_overlay_dicts()[0])
    for a, r in zip(results[0], ref):                       # and the reference's own cv2 calls on the same dict
        np.testing.assert_array_equal(a, r)
