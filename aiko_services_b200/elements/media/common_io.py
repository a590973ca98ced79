"""DataSource / DataTarget: same parameters and behaviour as R/elements/media/common_io.py
(``data_sources``, ``data_batch_size``, ``rate``; ``data_targets``), host-side logic only."""
import os
from pathlib import Path

from ...pipeline import PipelineElement
from ...stream import StreamEvent

__all__ = ["contains_all", "file_glob_difference", "DataSource", "DataTarget"]


def contains_all(source, match):
    return all(character in source for character in match)


def file_glob_difference(file_glob, filename):
    tokens = file_glob.split("*")
    start, end = tokens[0], (tokens[1] if len(tokens) > 1 else "")
    if filename.startswith(start) and filename.endswith(end):
        return filename[len(start):len(filename) - len(end)]
    return None


def _parse_list(text):
    """"(a b c)" or "a" -> ["a", "b", "c"] (the reference's S-expression list of URLs)."""
    text = str(text).strip()
    if text.startswith("(") and text.endswith(")"):
        text = text[1:-1]
    return text.split()


class DataSource(PipelineElement):
    def start_stream(self, stream, stream_id, use_create_frame=True):
        data_sources, found = self.get_parameter("data_sources")
        if not found:
            return StreamEvent.ERROR, {"diagnostic": 'Must provide "data_sources" parameter'}
        paths = []
        for data_source in _parse_list(data_sources):
            tokens = data_source.split("://")
            if len(tokens) == 1:
                path = tokens[0]
            else:
                if tokens[0] != "file":
                    return StreamEvent.ERROR, {"diagnostic": 'DataSource scheme must be "file://"'}
                path = tokens[1]
            file_glob = "*"
            if contains_all(path, "{}"):
                file_glob = os.path.basename(path).replace("{}", "*")
                path = os.path.dirname(path)
            path = Path(path)
            if not path.exists():
                return StreamEvent.ERROR, {"diagnostic": f'path "{path}" does not exist'}
            if path.is_file():
                paths.append((path, None))
            elif path.is_dir():
                for found_path in sorted(path.glob(file_glob)):
                    file_id = file_glob_difference(file_glob, found_path.name) if file_glob != "*" else None
                    paths.append((found_path, file_id))
            else:
                return StreamEvent.ERROR, {"diagnostic": f'"{path}" must be a file or a directory'}
        if use_create_frame and len(paths) == 1:
            self.create_frame(stream, {"paths": [paths[0][0]]})
        else:
            stream.variables["source_paths_generator"] = iter(paths)
            rate, _ = self.get_parameter("rate", default=None)
            self.create_frames(stream, self.frame_generator, rate=float(rate) if rate else None)
        return StreamEvent.OKAY, {}

    def frame_generator(self, stream, frame_id):
        data_batch_size, _ = self.get_parameter("data_batch_size", default=1)
        paths = []
        try:
            for _ in range(int(data_batch_size)):
                path, _file_id = next(stream.variables["source_paths_generator"])
                path = Path(path)
                if not path.is_file():
                    return StreamEvent.ERROR, {"diagnostic": f'path "{path}" must be a file'}
                paths.append(path)
        except StopIteration:
            pass
        if paths:
            return StreamEvent.OKAY, {"paths": paths}
        return StreamEvent.STOP, {"diagnostic": "All frames generated"}


class DataTarget(PipelineElement):
    def start_stream(self, stream, stream_id):
        data_targets, found = self.get_parameter("data_targets")
        if not found:
            return StreamEvent.ERROR, {"diagnostic": 'Must provide file "data_targets" parameter'}
        tokens = data_targets.split("://")
        if len(tokens) == 1:
            path = tokens[0]
        else:
            if tokens[0] != "file":
                return StreamEvent.ERROR, {"diagnostic": 'DataSource scheme must be "file://"'}
            path = tokens[1]
        stream.variables["target_file_id"] = 0
        stream.variables["target_path"] = path
        return StreamEvent.OKAY, {}
