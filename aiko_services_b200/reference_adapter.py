"""Run the GPU elements inside the UNMODIFIED reference runtime.

``install()`` needs the reference's ``aiko_services`` package to be importable.  It builds
element classes that derive from the reference's own ``aiko.PipelineElement`` (constructed the
way the reference constructs elements, R/elements/media/image_io.py:204-206) around the same
``logic_*`` functions this package's runtime uses, and registers them through the hook the
reference itself provides: ``load_module`` consults ``MODULES_LOADED`` before importing
(R/main/utilities/importer.py:22-26), so pre-seeding the entry
``"aiko_services.elements.media.image_io"`` (and ``.video_io``, ``.webcam_io``, ``.audio_io``) makes
every existing pipeline-definition JSON resolve ``ImageResize`` / ``ImageOverlay`` /
``VideoReadFile`` / ... to the GPU classes.  Nothing in the reference is edited.
"""
import types

MODULE_NAME = "aiko_services.elements.media.image_io"
AUDIO_MODULE_NAME = "aiko_services.elements.media.audio_io"
VIDEO_MODULE_NAME = "aiko_services.elements.media.video_io"
WEBCAM_MODULE_NAME = "aiko_services.elements.media.webcam_io"


def build_module():
    import aiko_services as aiko
    from aiko_services.elements.media import common_io as ref_common

    from .elements.media import image_io as gpu

    def make(name, protocol, logic, base):
        def __init__(self, context):
            context.set_protocol(f"{protocol}:0")
            context.get_implementation("PipelineElement").__init__(self, context)

        return type(name, (base,), {"__init__": __init__, "process_frame": logic,
                                    "__doc__": f"GPU {name} (aiko_services_b200) bound to the reference runtime"})

    module = types.ModuleType(MODULE_NAME + " [aiko_services_b200]")
    module.ImageResize = make("ImageResize", "image_resize", gpu.logic_image_resize, aiko.PipelineElement)
    module.ImageOverlay = make("ImageOverlay", "image_overlay", gpu.logic_image_overlay, aiko.PipelineElement)
    module.ImageConvert = make("ImageConvert", "image_convert", gpu.logic_image_convert, aiko.PipelineElement)
    module.ImageOutput = make("ImageOutput", "image_output",
                              lambda self, stream, images: (aiko.StreamEvent.OKAY, {"images": images}),
                              aiko.PipelineElement)
    module.ImageReadFile = make("ImageReadFile", "image_read_file", gpu.logic_image_read_file, ref_common.DataSource)
    module.ImageWriteFile = make("ImageWriteFile", "image_write_file", gpu.logic_image_write_file,
                                 ref_common.DataTarget)
    module.__all__ = ["ImageConvert", "ImageOutput", "ImageOverlay", "ImageReadFile", "ImageResize", "ImageWriteFile"]
    return module


def build_audio_module():
    """PE_FFT / PE_AudioResampler / PE_AudioFilter for ``aiko_services.elements.media.audio_io`` (in the
    reference they exist only as dead code inside a string literal, R/elements/media/audio_io.py:162-642,
    so a definition naming them cannot load there at all)."""
    import aiko_services as aiko
    from aiko_services.elements.media import audio_io as ref_audio

    from .elements.media import audio_io as gpu

    def make(name, protocol, cls):
        def __init__(self, context):
            context.set_protocol(f"{protocol}:0")
            context.get_implementation("PipelineElement").__init__(self, context)

        return type(name, (aiko.PipelineElement,), {"__init__": __init__, "process_frame": cls.process_frame,
                                                    "__doc__": cls.__doc__})

    module = types.ModuleType(AUDIO_MODULE_NAME + " [aiko_services_b200]")
    for name in getattr(ref_audio, "__all__", []):            # what the reference's module does export stays
        if hasattr(ref_audio, name):                          # (its __all__ names classes it never defines)
            setattr(module, name, getattr(ref_audio, name))
    module.PE_FFT = make("PE_FFT", "fft", gpu.PE_FFT)
    module.PE_AudioResampler = make("PE_AudioResampler", "audio_resample", gpu.PE_AudioResampler)
    module.PE_AudioFilter = make("PE_AudioFilter", "audio_filter", gpu.PE_AudioFilter)
    return module


def build_video_modules():
    """The reference's own VideoReadFile / VideoWriteFile / VideoReadWebcam with only their per-frame
    pixel work replaced (video_io.py:129,299; webcam_io.py:121-129): capture, codecs, file iteration
    and shared settings stay the reference's code."""
    import aiko_services as aiko
    from aiko_services.elements.media import video_io as ref_video
    from aiko_services.elements.media import webcam_io as ref_webcam
    import numpy as np

    from .elements.media.image_io import _apply

    class VideoReadFile(ref_video.VideoReadFile):
        def frame_generator(self, stream, frame_id):
            event, data = super().frame_generator(stream, frame_id)
            return event, data

        def video_frame_iterator(self, video_capture):
            stream, _ = self.get_stream()
            while True:
                status, image_bgr = video_capture.read()
                if not status:
                    break
                yield _apply(stream, [image_bgr], ("convert", "BGR2RGB"))[0]

    class VideoWriteFile(ref_video.VideoWriteFile):
        def process_frame(self, stream, images):
            if not stream.variables["video_writer"]:
                shape = images[0].shape
                stream.variables["video_writer"] = self._create_video_writer(
                    stream.variables["video_path"], (shape[1], shape[0]))
            for image in _apply(stream, list(images), ("convert", "RGB2BGR")):
                stream.variables["video_writer"].write(np.ascontiguousarray(np.asarray(image)))
            return aiko.StreamEvent.OKAY, {}

    class VideoReadWebcam(ref_webcam.VideoReadWebcam):
        def frame_generator(self, stream, frame_id):
            frame_data = None
            if self.video_capture and self.video_capture.isOpened():
                status, image_bgr = self.video_capture.read()
                if status:
                    if frame_id % 10 == 0:
                        self.ec_producer.update("frame_id", frame_id)
                    images = _apply(stream, [image_bgr], ("convert", "BGR2RGB" if self.share["color"] else "BGR2GRAY"))
                    flip = self.share["flip"]
                    if flip == "both" or flip == "horizontal":
                        images = _apply(stream, images, ("flip", 1))
                    if flip == "both" or flip == "vertical":
                        images = _apply(stream, images, ("flip", 0))
                    frame_data = {"images": images}
            return aiko.StreamEvent.OKAY, frame_data

    video = types.ModuleType(VIDEO_MODULE_NAME + " [aiko_services_b200]")
    for name in ref_video.__all__:
        if hasattr(ref_video, name):
            setattr(video, name, getattr(ref_video, name))
    video.VideoReadFile, video.VideoWriteFile = VideoReadFile, VideoWriteFile
    webcam = types.ModuleType(WEBCAM_MODULE_NAME + " [aiko_services_b200]")
    webcam.VideoReadWebcam = VideoReadWebcam
    return video, webcam


def install():
    """Make the reference's loader hand out the GPU elements for image_io, video_io, webcam_io and
    audio_io; returns the image module it will serve."""
    from aiko_services.main.utilities import importer
    module = build_module()
    importer.MODULES_LOADED[MODULE_NAME] = module
    importer.MODULES_LOADED[AUDIO_MODULE_NAME] = build_audio_module()
    video, webcam = build_video_modules()
    importer.MODULES_LOADED[VIDEO_MODULE_NAME] = video
    importer.MODULES_LOADED[WEBCAM_MODULE_NAME] = webcam
    return module


def uninstall():
    from aiko_services.main.utilities import importer
    for name in (MODULE_NAME, AUDIO_MODULE_NAME, VIDEO_MODULE_NAME, WEBCAM_MODULE_NAME):
        importer.MODULES_LOADED.pop(name, None)
