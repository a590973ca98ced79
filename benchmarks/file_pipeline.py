"""BASELINE config 1 as a user runs it: a pipeline definition ImageReadFile -> ImageResize -> ImageWriteFile over JPEG
files on disk (tmpfs), this package's runtime with decode / encode on the GPU, against the same definition with the
host codecs (Pillow) in the same runtime.  frames/s, wall clock, files written included."""
import os
import shutil
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from aiko_services_b200.pipeline import Pipeline  # noqa: E402


def definition(root, gpu, batch):
    mod = "aiko_services.elements.media.image_io"
    read = {"data_sources": f"(file://{root}/in_{{}}.jpg)", "data_batch_size": batch}
    write = {"data_targets": f"file://{root}/out_{{:04d}}.jpg"}
    if gpu:
        read["decode"] = "gpu"
        write["encode"] = "gpu"
    return {"version": 0, "name": "p_files", "runtime": "python", "graph": ["(ImageReadFile ImageResize ImageWriteFile)"],
            "parameters": {"resolution": "640x480"},
            "elements": [
                {"name": "ImageReadFile", "parameters": read, "input": [{"name": "paths", "type": "[Path]"}],
                 "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": mod}}},
                {"name": "ImageResize", "input": [{"name": "images", "type": "[image]"}],
                 "output": [{"name": "images", "type": "[image]"}], "deploy": {"local": {"module": mod}}},
                {"name": "ImageWriteFile", "parameters": write, "input": [{"name": "images", "type": "[image]"}], "output": [],
                 "deploy": {"local": {"module": mod}}}]}


def main():
    n_files = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    root = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        for k in range(n_files):
            with open(os.path.join(root, f"in_{k:04d}.jpg"), "wb") as f:
                f.write(bench.jpeg_blob(7000 + k % 32))
        if os.environ.get("PROFILE"):
            import cProfile
            import pstats
            p = Pipeline(definition(root, True, 32))
            assert p.create_stream("1")
            p.run()
            for f in os.listdir(root):
                if f.startswith("out_"):
                    os.remove(os.path.join(root, f))
            p = Pipeline(definition(root, True, 32))
            assert p.create_stream("1")
            prof = cProfile.Profile()
            prof.enable()
            p.run()
            prof.disable()
            pstats.Stats(prof).sort_stats("tottime").print_stats(16)
            return
        for gpu in (True, False, True):
            p = Pipeline(definition(root, gpu, 32))
            assert p.create_stream("1")
            t0 = time.perf_counter()
            p.run()
            dt = time.perf_counter() - t0
            written = len([f for f in os.listdir(root) if f.startswith("out_")])
            print(f"{'gpu codecs' if gpu else 'host codecs (Pillow)'}: {written} files, {written / dt:.0f} frames/s")
            for f in os.listdir(root):
                if f.startswith("out_"):
                    os.remove(os.path.join(root, f))
    finally:
        shutil.rmtree(root, ignore_errors=True)


if __name__ == "__main__":
    main()
