"""Run the GPU elements inside the UNMODIFIED reference runtime.

``install()`` needs the reference's ``aiko_services`` package to be importable.  It builds
element classes that derive from the reference's own ``aiko.PipelineElement`` (constructed the
way the reference constructs elements, R/elements/media/image_io.py:204-206) around the same
``logic_*`` functions this package's runtime uses, and registers them through the hook the
reference itself provides: ``load_module`` consults ``MODULES_LOADED`` before importing
(R/main/utilities/importer.py:22-26), so pre-seeding the entry
``"aiko_services.elements.media.image_io"`` makes every existing pipeline-definition JSON
resolve ``ImageResize`` / ``ImageOverlay`` / ... to the GPU classes.  Nothing in the reference
is edited.
"""
import types

MODULE_NAME = "aiko_services.elements.media.image_io"


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


def install():
    """Make the reference's loader hand out the GPU elements; returns the module it will serve."""
    from aiko_services.main.utilities import importer
    module = build_module()
    importer.MODULES_LOADED[MODULE_NAME] = module
    return module


def uninstall():
    from aiko_services.main.utilities import importer
    importer.MODULES_LOADED.pop(MODULE_NAME, None)
