"""Video elements with the reference's names, inputs, outputs and parameters
(R/elements/media/video_io.py).  Container demux and the video codec stay where the reference has
them -- OpenCV's ``cv2.VideoCapture`` / ``cv2.VideoWriter`` on the host (NVDEC / NVENC need headers
this image does not have) -- but the per-frame pixel work of those elements runs on the GPU:

  VideoReadFile   BGR -> RGB of every decoded frame (video_io.py:129)    -> aiko_b200_swap_rb_u8
  VideoWriteFile  RGB -> BGR before the encoder        (video_io.py:299)    -> fused into whatever
                  produced the frames (resize / overlay), or aiko_b200_swap_rb_u8
  VideoShow       windowing: unavailable on this box; the frames are dropped (headless sink)

Frames leave VideoReadFile as device-resident ``DeviceImage`` items, so a following ImageResize /
ImageOverlay never sees host pixels.  ``cv2`` is imported here for capture and codecs only.
"""
from pathlib import Path

import numpy as np

from ...device_frames import DeviceImage
from ...pipeline import PipelineElement
from ...stream import StreamEvent
from .common_io import DataSource, DataTarget, contains_all
from .image_io import _apply, stager_for

__all__ = ["VideoOutput", "VideoReadFile", "VideoSample", "VideoShow", "VideoWriteFile"]


def _cv2():
    try:
        import cv2
    except ModuleNotFoundError as error:       # the reference warns at import and fails at first use
        raise RuntimeError("video_io: OpenCV (cv2) is needed for video capture and codecs") from error
    return cv2


class VideoOutput(PipelineElement):
    def process_frame(self, stream, images):
        return StreamEvent.OKAY, {"images": images}


class VideoReadFile(DataSource):
    """video_io.py:114-177.  One decoded frame per pipeline frame, as in the reference; the optional
    parameter ``video_batch_size`` (default 1) puts that many consecutive frames of the file into one
    pipeline frame, which is what lets the GPU elements downstream work on batches."""

    def start_stream(self, stream, stream_id):
        stream.variables["video_capture"] = None
        stager_for(stream)               # created here, on the pipeline's thread, not by the generator thread
        return super().start_stream(stream, stream_id, use_create_frame=False)

    def _next_capture(self, stream):
        """Opens the next file of the stream; -> None when there is none left, or an error tuple."""
        cv2 = _cv2()
        try:
            item = next(stream.variables["source_paths_generator"])
        except StopIteration:
            return None
        if not item.path.is_file():
            return StreamEvent.ERROR, {"diagnostic": f'path "{item.path}" must be a file'}
        capture = cv2.VideoCapture(str(item.path))
        if not capture.isOpened():
            return StreamEvent.ERROR, {"diagnostic": f"Couldn't open video file: {item.path}"}
        stream.variables["video_capture"] = capture
        return capture

    def frame_generator(self, stream, frame_id):
        count, _ = self.get_parameter("video_batch_size", default=1)
        frames = []
        while len(frames) < max(int(count), 1):
            capture = stream.variables["video_capture"]
            if capture is None:
                capture = self._next_capture(stream)
                if capture is None:
                    break
                if isinstance(capture, tuple):
                    return capture
            status, image_bgr = capture.read()
            if status:
                frames.append(image_bgr)
            else:
                capture.release()
                stream.variables["video_capture"] = None
        if not frames:
            return StreamEvent.STOP, {"diagnostic": "End of video file(s)"}
        # BGR -> RGB (video_io.py:129) on the GPU: the frames enter the pipeline device-resident
        return StreamEvent.OKAY, {"images": _apply(stream, frames, ("convert", "BGR2RGB"))}

    def process_frame(self, stream, images):
        return StreamEvent.OKAY, {"images": images}

    def stop_stream(self, stream, stream_id):
        capture = stream.variables.get("video_capture")
        if capture is not None and capture.isOpened():
            capture.release()
        stream.variables["video_capture"] = None
        return StreamEvent.OKAY, {}


class VideoSample(PipelineElement):
    """video_io.py:181-193: keeps every ``sample_rate``-th frame."""

    def process_frame(self, stream, images):
        sample_rate, _ = self.get_parameter("sample_rate", 1)
        if stream.frame_id % int(sample_rate):
            return StreamEvent.DROP_FRAME, {}
        return StreamEvent.OKAY, {"images": images}


class VideoShow(PipelineElement):
    """video_io.py:198-237 opens an OpenCV window: unavailable on this box (no display, no graphics
    stack).  Kept so that definitions naming it load; the frames end here unseen."""
    _told = False

    def process_frame(self, stream, images):
        if not VideoShow._told:
            VideoShow._told = True
            self.logger.warning("VideoShow: no display on this machine; frames are not shown")
        return StreamEvent.OKAY, {}


class VideoWriteFile(DataTarget):
    """video_io.py:246-308: parameters ``format`` (fourcc, default MP4V), ``frame_rate`` (30.0),
    ``resolution`` (default: the first frame's)."""

    def start_stream(self, stream, stream_id):
        event, diagnostic = super().start_stream(stream, stream_id)
        if event == StreamEvent.OKAY:
            path = stream.variables["target_path"]
            if contains_all(path, "{}"):
                path = path.format(stream.variables["target_file_id"])
            stream.variables["video_path"] = path
            stream.variables["video_writer"] = None
            diagnostic = {}
        return event, diagnostic

    def _create_video_writer(self, path, resolution, format="MP4V", frame_rate=30.0):
        cv2 = _cv2()
        format, _ = self.get_parameter("format", format)
        frame_rate, _ = self.get_parameter("frame_rate", frame_rate)
        resolution, _ = self.get_parameter("resolution", resolution)
        if isinstance(resolution, str):
            width, height = resolution.split("x")
            resolution = (int(width), int(height))
        Path(path).parent.mkdir(exist_ok=True, parents=True)
        return cv2.VideoWriter(str(path), cv2.VideoWriter_fourcc(*format), float(frame_rate), tuple(resolution))

    def process_frame(self, stream, images):
        if not images:
            return StreamEvent.OKAY, {}
        writer = stream.variables["video_writer"]
        if writer is None:
            shape = images[0].shape
            writer = self._create_video_writer(stream.variables["video_path"], (shape[1], shape[0]))
            stream.variables["video_writer"] = writer
        # RGB -> BGR (video_io.py:299) joins the frames' lazy recipe: fused into the resize / overlay
        # kernel when one produced them, one swap kernel otherwise; then one copy to the host
        for image in _apply(stream, list(images), ("convert", "RGB2BGR")):
            writer.write(np.ascontiguousarray(np.asarray(image)))
        return StreamEvent.OKAY, {}

    def stop_stream(self, stream, stream_id):
        writer = stream.variables.get("video_writer")
        if writer is not None:
            writer.release()
            stream.variables["video_writer"] = None
        return StreamEvent.OKAY, {}
