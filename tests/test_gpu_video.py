"""GPU: the video / webcam elements (capture and codec on the host through OpenCV, as in the reference;
BGR<->RGB, BGR->GRAY and flips on the GPU) and the torch.ops.aiko_b200 custom ops.  Expected values are
computed with the reference elements' own cv2 calls on the frames cv2 decodes from the same file."""
import numpy as np
import pytest
import torch

from aiko_services_b200.pipeline import Pipeline
from golden import cases
from oracle import image as O

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")

VIDEO_IO = "aiko_services.elements.media.video_io"
WEBCAM_IO = "aiko_services.elements.media.webcam_io"
IMAGE_IO = "aiko_services.elements.media.image_io"


def write_video(path, n=7, h=120, w=160):
    writer = cv2.VideoWriter(str(path), cv2.VideoWriter_fourcc(*"MJPG"), 30.0, (w, h))
    assert writer.isOpened(), "OpenCV cannot write MJPG/AVI here"
    for k in range(n):
        writer.write(cases.frame(7000 + k, h, w, 3))
    writer.release()


def decoded_bgr(path):
    capture = cv2.VideoCapture(str(path))
    frames = []
    while True:
        ok, f = capture.read()
        if not ok:
            break
        frames.append(f)
    capture.release()
    return frames


def element(name, module, inputs, outputs, parameters=None):
    e = {"name": name, "input": [{"name": k, "type": "any"} for k in inputs],
         "output": [{"name": k, "type": "any"} for k in outputs], "deploy": {"local": {"module": module}}}
    if parameters:
        e["parameters"] = parameters
    return e


def test_video_read_resize_write(cuda, tmp_path):
    """VideoReadFile -> ImageResize -> VideoWriteFile, definition as the reference's video pipelines spell it."""
    import queue
    write_video(tmp_path / "in.avi")
    q = queue.Queue()
    d = {"version": 0, "name": "p_video", "runtime": "python",
         "graph": ["(VideoReadFile ImageResize VideoOutput VideoWriteFile)"], "parameters": {"resolution": "80x60"},
         "elements": [
             element("VideoReadFile", VIDEO_IO, ["images"], ["images"], {"data_sources": f"(file://{tmp_path}/in.avi)"}),
             element("ImageResize", IMAGE_IO, ["images"], ["images"]),
             element("VideoOutput", VIDEO_IO, ["images"], ["images"]),
             element("VideoWriteFile", VIDEO_IO, ["images"], [], {"data_targets": f"file://{tmp_path}/out.avi", "format": "MJPG"})]}
    p = Pipeline(d)
    assert p.create_stream("v", queue_response=q)
    p.run()
    source = decoded_bgr(tmp_path / "in.avi")
    assert len(source) == 7
    got = []
    while not q.empty():
        got.append(q.get()[1]["images"])
    assert len(got) == 7
    expected_written = []
    for k, images in enumerate(got):
        rgb = cv2.cvtColor(source[k], cv2.COLOR_BGR2RGB)                 # video_io.py:129
        want = cv2.resize(rgb, (80, 60))                                 # image_io.py:216
        np.testing.assert_array_equal(np.asarray(images[0]), want)
        expected_written.append(cv2.cvtColor(want, cv2.COLOR_RGB2BGR))   # video_io.py:299
    # the file holds what cv2's encoder makes of exactly those BGR frames
    ref_path = tmp_path / "ref.avi"
    writer = cv2.VideoWriter(str(ref_path), cv2.VideoWriter_fourcc(*"MJPG"), 30.0, (80, 60))
    for f in expected_written:
        writer.write(f)
    writer.release()
    ours, theirs = decoded_bgr(tmp_path / "out.avi"), decoded_bgr(ref_path)
    assert len(ours) == len(theirs) == 7
    for a, b in zip(ours, theirs):
        np.testing.assert_array_equal(a, b)


def test_video_batches_and_sampling(cuda, tmp_path):
    write_video(tmp_path / "clip_0.avi", n=5)
    write_video(tmp_path / "clip_1.avi", n=4)
    import queue
    q = queue.Queue()
    d = {"version": 0, "name": "p_video2", "runtime": "python", "graph": ["(VideoReadFile VideoSample VideoOutput)"],
         "elements": [
             element("VideoReadFile", VIDEO_IO, ["images"], ["images"],
                     {"data_sources": f"(file://{tmp_path}/clip_{{}}.avi)", "video_batch_size": 3}),
             element("VideoSample", VIDEO_IO, ["images"], ["images"], {"sample_rate": 2}),
             element("VideoOutput", VIDEO_IO, ["images"], ["images"])]}
    p = Pipeline(d)
    assert p.create_stream("v", queue_response=q)
    p.run()
    batches = []
    while not q.empty():
        info, swag = q.get()
        batches.append((int(info["frame_id"]), len(swag["images"])))
    # 9 frames in batches of 3 (files are read back to back) -> pipeline frames 0, 1, 2; every second is kept
    assert batches == [(0, 3), (2, 3)]


@pytest.mark.parametrize("color,flip", [(True, "none"), (False, "horizontal"), (True, "both"), (True, "vertical")])
def test_webcam_colour_and_flip(cuda, tmp_path, color, flip):
    """VideoReadWebcam with a video file standing in for the camera: webcam_io.py:121-129 on the GPU."""
    import queue
    write_video(tmp_path / "cam.avi", n=3, h=96, w=128)
    q = queue.Queue()
    d = {"version": 0, "name": "p_cam", "runtime": "python", "graph": ["(VideoReadWebcam VideoOutput)"],
         "elements": [element("VideoReadWebcam", WEBCAM_IO, ["images"], ["images"],
                              {"path": str(tmp_path / "cam.avi"), "color": color, "flip": flip}),
                      element("VideoOutput", VIDEO_IO, ["images"], ["images"])]}
    p = Pipeline(d)
    assert p.create_stream("c", queue_response=q)
    p.run()
    source = decoded_bgr(tmp_path / "cam.avi")
    got = []
    while not q.empty():
        got.append(np.asarray(q.get()[1]["images"][0]))
    assert len(got) == 3
    for k in range(3):
        want = cv2.cvtColor(source[k], cv2.COLOR_BGR2RGB if color else cv2.COLOR_BGR2GRAY)
        if flip in ("both", "horizontal"):
            want = cv2.flip(want, 1)
        if flip in ("both", "vertical"):
            want = cv2.flip(want, 0)
        np.testing.assert_array_equal(got[k], want)


def test_torch_custom_ops(cuda):
    """torch.ops.aiko_b200.*: the C ABI as PyTorch custom ops (north_star), called directly."""
    from aiko_services_b200 import ops, torch_ops  # noqa: F401  (registers the ops)
    T = torch.ops.aiko_b200
    stream = torch.cuda.current_stream().cuda_stream
    frames = np.stack([cases.frame(8100 + k, 90, 150, 3) for k in range(3)])
    src = torch.from_numpy(frames).to(cuda)
    out = torch.empty((3, 45, 60, 3), dtype=torch.uint8, device=cuda)
    T.resize_linear_u8(src, out, True, 0, stream)
    for k in range(3):
        np.testing.assert_array_equal(out[k].cpu().numpy(), O.resize_linear(frames[k], 60, 45)[:, :, ::-1])
    gray = torch.empty((3, 90, 150), dtype=torch.uint8, device=cuda)
    T.bgr2gray_u8(src, gray, False, stream)
    np.testing.assert_array_equal(gray[1].cpu().numpy(), O.bgr2gray(frames[1]))
    flipped = torch.empty_like(src)
    T.flip_u8(src, flipped, 1, stream)
    np.testing.assert_array_equal(flipped[2].cpu().numpy(), frames[2][:, ::-1])
    swapped = torch.empty_like(src)
    T.swap_rb_u8(src, swapped, stream)
    np.testing.assert_array_equal(swapped[0].cpu().numpy(), frames[0][:, :, ::-1])
    ws = ops.bicubic_workspace(cuda, 90, 150, 45, 60)
    T.resize_bicubic_aa_u8(src, out, ws, stream)
    np.testing.assert_array_equal(out[0].cpu().numpy(), O.resize_bicubic_aa(frames[0], 60, 45))
    with pytest.raises(Exception):
        T.flip_u8(src, src, 1, stream)           # aliased pointers: the library's error, not a crash
    # JPEG files of the frames (image.save's bytes), straight through the op
    from aiko_services_b200 import _native
    capacity = 90 * 150 * 3
    nbytes = _native.lib().aiko_b200_jpeg_gpu_encode_workspace_bytes(3, 150, 90, 3, 2, 2, capacity)
    workspace = torch.empty(nbytes, dtype=torch.uint8, device=cuda)
    files = torch.empty((3, 2 * capacity), dtype=torch.uint8, device=cuda)
    sizes = torch.empty(3, dtype=torch.int32, device=cuda)
    T.jpeg_encode_u8(src, 75, 2, 2, workspace, capacity, files, sizes, stream)
    for k in range(3):
        assert files[k, :int(sizes[k])].cpu().numpy().tobytes() == O.jpeg_encode(frames[k])
