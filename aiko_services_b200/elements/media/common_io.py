"""Sources and targets of file-backed streams for this package's runtime.

What pipeline definitions and other elements rely on is kept as the reference has it
(R/elements/media/common_io.py:52-151): the parameters ``data_sources`` (a URL or an S-expression
list of URLs, ``file://`` or bare paths, ``{}`` standing for the varying part of a file name),
``data_batch_size`` (items per frame -- the knob BASELINE config 2 uses for its 32-frame batches),
``rate`` and ``data_targets``; frames carry ``{"paths": [Path, ...]}``; a target's location and
running file number live in ``stream.variables["target_path" / "target_file_id"]``; and the
diagnostics name the same conditions.  How that is computed is this module's own: a locator is
parsed once into a ``Locator``, expanded into an ordered list of ``Item``s, and a ``Batcher`` deals
them out -- three small pieces that the video source reuses for its own iteration.

(Inside the reference's own runtime these classes are not used at all: reference_adapter binds the
GPU elements to the reference's own DataSource / DataTarget.)
"""
import re
from dataclasses import dataclass
from pathlib import Path
from typing import List, Optional

from ...pipeline import PipelineElement
from ...stream import StreamEvent

__all__ = ["contains_all", "file_glob_difference", "Batcher", "DataSource", "DataTarget", "Item", "Locator",
           "LocatorError"]


def contains_all(source, match):
    """True when every character of ``match`` occurs in ``source`` (the reference's helper, used by
    its elements to spot the "{}" placeholder)."""
    return set(match) <= set(source)


def file_glob_difference(file_glob, filename):
    """The part of ``filename`` matched by the single ``*`` of ``file_glob``, or None."""
    head, star, tail = file_glob.partition("*")
    if not star:
        return None
    match = re.fullmatch(re.escape(head) + "(.*)" + re.escape(tail), filename)
    return match.group(1) if match else None


class LocatorError(ValueError):
    """A ``data_sources`` / ``data_targets`` value that cannot be used; str() is the diagnostic."""


@dataclass(frozen=True)
class Item:
    path: Path
    file_id: Optional[str] = None      # what "{}" stood for in this file's name


@dataclass(frozen=True)
class Locator:
    """One URL of ``data_sources`` / ``data_targets``: where, and which names under it."""
    location: str
    pattern: Optional[str] = None      # file-name pattern with "*" where the URL had "{}"

    @staticmethod
    def parse(url):
        scheme, separator, rest = str(url).partition("://")
        if separator and scheme != "file":
            raise LocatorError('DataSource scheme must be "file://"')
        location = rest if separator else scheme
        if "{" in location and "}" in location:
            folder, _, name = location.rpartition("/")
            return Locator(folder, name.replace("{}", "*"))
        return Locator(location)

    @staticmethod
    def parse_list(text):
        """``"(url url ...)"`` or a single url -> [Locator]"""
        return [Locator.parse(word) for word in str(text).strip().strip("()").split()]

    def expand(self) -> List[Item]:
        where = Path(self.location)
        if not where.exists():
            raise LocatorError(f'path "{where}" does not exist')
        if where.is_file():
            return [Item(where)]
        if not where.is_dir():
            raise LocatorError(f'"{where}" must be a file or a directory')
        pattern = self.pattern or "*"
        return [Item(found, file_glob_difference(pattern, found.name) if self.pattern else None)
                for found in sorted(where.glob(pattern))]


class Batcher:
    """Deals the items of a stream out ``size`` at a time, in order."""

    def __init__(self, items):
        self._items = iter(items)

    def __iter__(self):
        return self

    def __next__(self):              # one item (the video source walks files one by one)
        return next(self._items)

    def take(self, size):
        batch = []
        for item in self._items:
            batch.append(item)
            if len(batch) >= size:
                break
        return batch


class DataSource(PipelineElement):
    def start_stream(self, stream, stream_id, use_create_frame=True):
        data_sources, found = self.get_parameter("data_sources")
        if not found:
            return StreamEvent.ERROR, {"diagnostic": 'Must provide "data_sources" parameter'}
        try:
            items = [item for locator in Locator.parse_list(data_sources) for item in locator.expand()]
        except LocatorError as error:
            return StreamEvent.ERROR, {"diagnostic": str(error)}
        if use_create_frame and len(items) == 1:
            # a single item needs no generator thread
            self.create_frame(stream, {"paths": [items[0].path]})
            return StreamEvent.OKAY, {}
        stream.variables["source_paths_generator"] = Batcher(items)
        rate, _ = self.get_parameter("rate", default=None)
        self.create_frames(stream, self.frame_generator, rate=float(rate) if rate else None)
        return StreamEvent.OKAY, {}

    def frame_generator(self, stream, frame_id):
        size, _ = self.get_parameter("data_batch_size", default=1)
        batch = stream.variables["source_paths_generator"].take(max(int(size), 1))
        if not batch:
            return StreamEvent.STOP, {"diagnostic": "All frames generated"}
        for item in batch:
            if not item.path.is_file():
                return StreamEvent.ERROR, {"diagnostic": f'path "{item.path}" must be a file'}
        return StreamEvent.OKAY, {"paths": [item.path for item in batch]}


class DataTarget(PipelineElement):
    def start_stream(self, stream, stream_id):
        data_targets, found = self.get_parameter("data_targets")
        if not found:
            return StreamEvent.ERROR, {"diagnostic": 'Must provide file "data_targets" parameter'}
        scheme, separator, rest = str(data_targets).partition("://")
        if separator and scheme != "file":
            return StreamEvent.ERROR, {"diagnostic": 'DataSource scheme must be "file://"'}
        stream.variables["target_file_id"] = 0
        stream.variables["target_path"] = rest if separator else scheme
        return StreamEvent.OKAY, {}
