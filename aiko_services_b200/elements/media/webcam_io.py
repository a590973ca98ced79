"""VideoReadWebcam with the reference's name, parameters and shared settings
(R/elements/media/webcam_io.py:61-144).  The camera (or any path cv2.VideoCapture opens) is read on
the host; what the reference then does to every frame on the CPU runs on the GPU:

  share["color"] true   BGR -> RGB   (webcam_io.py:122)  -> aiko_b200_swap_rb_u8
  share["color"] false  BGR -> GRAY  (webcam_io.py:124)  -> aiko_b200_bgr2gray_u8
  share["flip"]  horizontal / vertical / both  (webcam_io.py:125-129)  -> aiko_b200_flip_u8

``cv2`` is imported here for the capture device only.
"""
from ...stream import StreamEvent
from .common_io import DataSource
from .image_io import _apply, stager_for

__all__ = ["VideoReadWebcam"]

_DEFAULT_CAMERA_PATHNAME = 0
_FLIP_CODES = {"horizontal": (1,), "vertical": (0,), "both": (1, 0)}     # cv2.flip codes, in the reference's order


class VideoReadWebcam(DataSource):
    def __init__(self, context):
        super().__init__(context)
        self.path_current = None
        self.stream_started = 0
        self.video_capture = None
        self.share.setdefault("color", True)
        self.share.setdefault("flip", "none")       # "both", "horizontal", "vertical"
        self.share["frame_id"] = -1
        self.share.setdefault("path", _DEFAULT_CAMERA_PATHNAME)

    def _close_camera(self):
        if self.video_capture is not None:
            self.video_capture.release()
        self.video_capture = None
        self.path_current = None

    def _open_camera(self, path):
        import cv2
        if self.stream_started > 0:
            if self.video_capture:
                self._close_camera()
            if isinstance(path, str) and path.isdigit():
                path = int(path)
            self.video_capture = cv2.VideoCapture(path)
            if self.video_capture.isOpened():
                self.path_current = path
                self.share["path"] = path
            else:
                self.logger.error(f"Open camera: {path} failed")
                self.video_capture = None

    def start_stream(self, stream, stream_id):
        self.stream_started += 1
        path, _ = self.get_parameter("path", _DEFAULT_CAMERA_PATHNAME)
        stager_for(stream)
        self._open_camera(path)
        self.create_frames(stream, self.frame_generator, rate=None)
        return StreamEvent.OKAY, {}

    def frame_generator(self, stream, frame_id):
        frame_data = None
        if self.video_capture and self.video_capture.isOpened():
            status, image_bgr = self.video_capture.read()
            if status:
                if frame_id % 10 == 0:
                    self.share["frame_id"] = frame_id
                color = self.share["color"]
                if isinstance(color, str):
                    color = color.lower() == "true"
                images = _apply(stream, [image_bgr], ("convert", "BGR2RGB" if color else "BGR2GRAY"))
                for code in _FLIP_CODES.get(self.share["flip"], ()):
                    images = _apply(stream, images, ("flip", code))
                frame_data = {"images": images}
            elif not isinstance(self.path_current, int):
                # a file stands in for the camera (tests, replays): its end ends the stream
                return StreamEvent.STOP, {"diagnostic": "Camera stopped"}
        return StreamEvent.OKAY, frame_data

    def process_frame(self, stream, images):
        return StreamEvent.OKAY, {"images": images}

    def stop_stream(self, stream, stream_id):
        self.stream_started -= 1
        self._close_camera()
        return StreamEvent.OKAY, {}
