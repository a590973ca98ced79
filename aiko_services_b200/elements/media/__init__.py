"""GPU implementations of the reference's media elements (R/elements/media/), same class names."""
from .common_io import DataSource, DataTarget, contains_all, file_glob_difference  # noqa: F401
from .image_io import (ImageConvert, ImageOutput, ImageOverlay, ImageReadFile, ImageResize,  # noqa: F401
                       ImageToTensor, ImageWriteFile)
from .video_io import VideoOutput, VideoReadFile, VideoSample, VideoShow, VideoWriteFile  # noqa: F401
from .webcam_io import VideoReadWebcam  # noqa: F401
