"""A small pipeline runtime for the media hot path, API-compatible with the subset of the
reference's R/main/pipeline.py that the path needs, plus batching across streams.

What is mirrored (same names, argument meaning and error behaviour):
  * the pipeline-definition JSON: "graph" S-expressions incl. ``(from: to)`` renames, "parameters",
    "elements" with "input" / "output" / "parameters" / "deploy.local.module|class_name"
    (R/main/pipeline.py:896-973, R/main/utilities/graph.py:115-152);
  * ``PipelineElement``: ``process_frame(stream, **inputs) -> (StreamEvent, dict)``,
    ``start_stream`` / ``stop_stream``, ``get_parameter`` with the reference's resolution order
    (R/main/pipeline.py:422-456), ``create_frame(s)``, ``get_stream``, ``my_id``, ``logger``;
  * ``Pipeline.create_stream / process_frame / destroy_stream`` and ``queue_response``
    (R/main/pipeline.py:725-860, 1018-1113): the per-frame graph walk, ``frame.swag``,
    ERROR / STOP / DROP_FRAME handling, per-element wall-clock metrics.

What is not here (out of scope): MQTT, actors and mailboxes, the registrar,
remote elements, leases, the dashboard.  A pipeline is driven by calling ``process_frame`` /
``process_frames`` directly or by elements' ``create_frame(s)`` from ``start_stream``.

What is new: ``process_frames`` takes frames of MANY streams at once.  Elements that declare
``batch_across_streams = True`` receive them in one ``process_frame_batch`` call (frames keyed
by stream_id), so one kernel launch serves every concurrent stream; other elements are called
frame by frame exactly as in the reference.
"""
import importlib
import importlib.util
import json
import logging
import os
import queue
import re
import threading
import time
import traceback
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Any, Dict, List, Tuple

from .metrics import GpuSeconds as _GpuSeconds, set_element_tag as _set_element_tag
from .stream import DEFAULT_STREAM_ID, FIRST_FRAME_ID, Frame, Stream, StreamEvent, StreamEventName, StreamState

__all__ = ["Pipeline", "PipelineDefinition", "PipelineElement", "PipelineElementDefinition", "parse_graph"]

_LOGGER = logging.getLogger("aiko_b200.pipeline")


# --------------------------------------------------------------------------- definition
@dataclass
class PipelineElementDefinition:
    name: str
    input: List[Dict[str, str]]
    output: List[Dict[str, str]]
    parameters: Dict[str, Any]
    deploy: Dict[str, Any]


@dataclass
class PipelineDefinition:
    version: int
    name: str
    runtime: str
    graph: List[str]
    parameters: Dict[str, Any]
    elements: List[PipelineElementDefinition]
    map_in_nodes: Dict[str, Any] = field(default_factory=dict)
    map_out_nodes: Dict[str, Any] = field(default_factory=dict)


def _tokenize(text):
    return re.findall(r"\(|\)|[^\s()]+", text)


def _parse_sexpr(tokens, pos=0):
    """-> (nested list, next position).  ``(a: b)`` forms become {"a": "b"} dictionaries."""
    assert tokens[pos] == "("
    pos += 1
    items = []
    while tokens[pos] != ")":
        if tokens[pos] == "(":
            sub, pos = _parse_sexpr(tokens, pos)
            items.append(sub)
        else:
            items.append(tokens[pos])
            pos += 1
    pos += 1
    if len(items) >= 2 and isinstance(items[0], str) and items[0].endswith(":"):
        d, i = {}, 0
        while i + 1 < len(items) and isinstance(items[i], str) and items[i].endswith(":"):
            d[items[i][:-1]] = items[i + 1]
            i += 2
        return d, pos
    return items, pos


def parse_graph(graph_definition):
    """Graph S-expressions -> (heads, successors, renames).

    ``"(A B C)"`` is a chain; ``"(A (B D) C)"`` gives A successors B and C, and B successor D;
    ``"(A B (x: y) C)"`` maps A's output ``x`` to B's input ``y``
    (R/main/utilities/graph.py:115-152, R/main/pipeline.py:573-582).
    """
    heads, successors, renames = OrderedDict(), OrderedDict(), []

    def add(node, succ):
        successors.setdefault(node, OrderedDict())
        if succ is not None:
            successors[node][succ] = succ
            successors.setdefault(succ, OrderedDict())

    def walk(node, rest):
        last = None
        for item in rest:
            if isinstance(item, dict):
                if last is not None:
                    renames.append((node, last, item))
            elif isinstance(item, list):
                add(node, item[0])
                last = item[0]
                walk(item[0], item[1:])
            else:
                add(node, item)
                last = item

    for text in graph_definition:
        tree, _ = _parse_sexpr(_tokenize(text))
        head = tree[0]
        heads[head] = head
        add(head, None)
        # "(A B C)" in the reference means A -> B and A -> C in declaration order, which for a
        # linear list executes as the chain A, B, C (get_path is a pre-order walk).
        walk(head, tree[1:])
    return heads, successors, renames


def execution_order(heads, successors, head_name=None):
    """Pre-order walk where a node reached again moves to the later position (graph.py:61-78)."""
    order = OrderedDict()

    def visit(node):
        if node in order:
            del order[node]
        order[node] = None
        for s in successors.get(node, {}):
            visit(s)

    if heads:
        visit(head_name if head_name in heads else next(iter(heads)))
    return list(order)


def parse_pipeline_definition(pathname_or_dict):
    if isinstance(pathname_or_dict, dict):
        d = pathname_or_dict
    else:
        with open(pathname_or_dict) as f:
            d = json.load(f)
    for key in ("version", "name", "runtime", "graph", "elements"):
        if key not in d:
            raise SystemExit(f"PipelineDefinition: missing field {key!r}")
    if d["version"] != 0:
        raise SystemExit(f"PipelineDefinition: version must be 0, but is {d['version']}")
    if d["runtime"] != "python":
        raise SystemExit(f'PipelineDefinition: runtime must be "python", but is "{d["runtime"]}"')
    elements = []
    for e in d["elements"]:
        deploy = e.get("deploy", {})
        if "local" not in deploy:
            raise SystemExit(f"PipelineDefinition: element {e.get('name')}: only deploy.local is supported here")
        elements.append(PipelineElementDefinition(e["name"], e.get("input", []), e.get("output", []),
                                                  e.get("parameters", {}), deploy))
    definition = PipelineDefinition(d["version"], d["name"], d["runtime"], d["graph"], d.get("parameters", {}),
                                    elements)
    _, _, renames = parse_graph(definition.graph)
    for predecessor, successor, mapping in renames:
        definition.map_out_nodes.setdefault(predecessor, {})[successor] = mapping
        definition.map_in_nodes.setdefault(successor, {})[predecessor] = mapping
    return definition


def _load_module(descriptor):
    if descriptor.endswith(".py"):
        spec = importlib.util.spec_from_file_location("aiko_b200_user_module", descriptor)
        module = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(module)
        return module
    # existing definitions name the reference's modules: serve them from this package
    if descriptor.startswith("aiko_services.elements."):
        descriptor = "aiko_services_b200.elements." + descriptor[len("aiko_services.elements."):]
    return importlib.import_module(descriptor)


# --------------------------------------------------------------------------- element
class PipelineElement:
    """Base class of elements run by this runtime (the role of the reference's
    PipelineElementImpl, R/main/pipeline.py:337-480)."""

    batch_across_streams = False

    def __init__(self, context):
        # ``context`` mirrors what the reference hands to an element constructor
        self.name = context["name"]
        self.definition = context["definition"]
        self.pipeline = context["pipeline"]
        self.logger = logging.getLogger(f"aiko_b200.{self.name}")
        self.share = dict(self.definition.parameters)

    # -- frames
    def create_frame(self, stream, frame_data, frame_id=None):
        frame_id = frame_id if frame_id else stream.frame_id
        self.pipeline.create_frame(stream, frame_data, frame_id)

    def create_frames(self, stream, frame_generator, frame_id=FIRST_FRAME_ID, rate=None):
        with self.pipeline._lock:
            self.pipeline._generators += 1
        threading.Thread(target=self.pipeline._frames_generator, daemon=True,
                         args=(self, stream, frame_generator, int(frame_id), rate)).start()

    # -- parameters: stream "Element.name" -> element definition (live share wins) -> stream "name"
    #    -> pipeline definition -> default  (R/main/pipeline.py:422-456)
    def get_parameter(self, name, default=None, use_pipeline=True, self_share_priority=True):
        value, found = None, False
        stream_parameters = self._get_stream_parameters()
        element_name = f"{self.definition.name}.{name}"
        if element_name in stream_parameters:
            value, found = stream_parameters[element_name], True
        elif name in self.definition.parameters:
            value = self.share[name] if (self_share_priority and name in self.share) else self.definition.parameters[name]
            found = True
        if not found and use_pipeline:
            if name in stream_parameters:
                value, found = stream_parameters[name], True
            elif name in self.pipeline.definition.parameters:
                value, found = self.pipeline.definition.parameters[name], True
        if not found and default is not None:
            value = default
        return value, found

    def _get_stream_parameters(self):
        try:
            stream, _ = self.get_stream()
            return stream.parameters if stream else {}
        except (AttributeError, AssertionError):
            return {}

    def get_stream(self):
        return self.pipeline.get_stream()

    def my_id(self, all=False):
        stream, frame_id = self.get_stream()
        return f"{self.name if all else ''}<{stream.stream_id}:{frame_id}>"

    def process_frame(self, stream, **kwargs) -> Tuple[int, dict]:
        raise NotImplementedError

    def start_stream(self, stream, stream_id):
        return StreamEvent.OKAY, None

    def stop_stream(self, stream, stream_id):
        return StreamEvent.OKAY, None


# --------------------------------------------------------------------------- pipeline
class Pipeline:
    def __init__(self, definition, stream_parameters=None):
        self.definition = definition if isinstance(definition, PipelineDefinition) \
            else parse_pipeline_definition(definition)
        self.logger = _LOGGER
        self.thread_local = threading.local()
        self.streams: Dict[str, Stream] = OrderedDict()
        self._pending = queue.Queue()           # (stream_id, frame_id, frame_data) from create_frame(s)
        self._generators = 0                    # frame-generator threads still running
        self._lock = threading.RLock()
        heads, successors, _ = parse_graph(self.definition.graph)
        self._heads, self._successors = heads, successors
        by_name = {e.name: e for e in self.definition.elements}
        self.elements: Dict[str, PipelineElement] = OrderedDict()
        for node in successors:
            if node not in by_name:
                raise SystemExit(f"PipelineDefinition: graph node {node} has no element definition")
            edef = by_name[node]
            local = edef.deploy["local"]
            module = _load_module(local["module"])
            cls = getattr(module, local.get("class_name", edef.name), None)
            if cls is None:
                raise SystemExit(f"PipelineDefinition: PipelineElement {edef.name}: "
                                 f"Module {local['module']} could not be loaded")
            self.elements[node] = cls({"name": node, "definition": edef, "pipeline": self})

    # -- thread-local "current stream" (R/main/pipeline.py:595-610)
    def _enable_thread_local(self, stream, frame_id):
        self.thread_local.stream, self.thread_local.frame_id = stream, frame_id

    def _disable_thread_local(self):
        self.thread_local.stream, self.thread_local.frame_id = None, None

    def get_stream(self):
        stream = getattr(self.thread_local, "stream", None)
        assert stream, "self.thread_local.stream must be assigned"
        return stream, self.thread_local.frame_id

    def graph_path(self, head=None):
        return execution_order(self._heads, self._successors, head)

    # -- streams
    def create_stream(self, stream_id, graph_path=None, parameters=None, grace_time=None, queue_response=None,
                      topic_response=None):
        stream_id = str(stream_id)
        with self._lock:
            if stream_id in self.streams:
                self.logger.error(f"Create stream: {stream_id} already exists")
                return False
            stream = Stream(stream_id=stream_id, graph_path=graph_path, parameters=dict(parameters or {}),
                            queue_response=queue_response, topic_response=topic_response)
            self.streams[stream_id] = stream
        self._enable_thread_local(stream, FIRST_FRAME_ID)
        try:
            for node in self.graph_path(graph_path):
                try:
                    event, diagnostic = self.elements[node].start_stream(stream, stream_id)
                except Exception:
                    event, diagnostic = StreamEvent.ERROR, {"diagnostic": traceback.format_exc()}
                state = self._process_stream_event(node, stream, event, diagnostic or {})
                if state != StreamState.RUN:    # never overwrite a STOP already set by a fast frame generator
                    stream.state = state
                if stream.state == StreamState.ERROR:
                    return False
        finally:
            self._disable_thread_local()
        return True

    def destroy_stream(self, stream_id, graceful=False):
        stream_id = str(stream_id)
        with self._lock:
            stream = self.streams.pop(stream_id, None)
        if stream is None:
            return False
        self._enable_thread_local(stream, stream.frame_id)
        try:
            for node in self.graph_path(stream.graph_path):
                try:
                    self.elements[node].stop_stream(stream, stream_id)
                except Exception:
                    self.logger.error(f"stop_stream(): {node}: {traceback.format_exc()}")
        finally:
            self._disable_thread_local()
        stream.state = StreamState.STOP if stream.state == StreamState.RUN else stream.state
        return True

    def _process_stream_event(self, element_name, stream, event, diagnostic):
        state = StreamState.RUN
        if event == StreamEvent.DROP_FRAME:
            state = StreamState.DROP_FRAME
        elif event == StreamEvent.STOP:
            state = StreamState.STOP
        elif event == StreamEvent.ERROR:
            state = StreamState.ERROR
            text = diagnostic.get("diagnostic", "No diagnostic provided") if isinstance(diagnostic, dict) else diagnostic
            self.logger.error(f"{element_name.upper()}: {StreamEventName[event]} stream <{stream.stream_id}> {text}")
        return state

    # -- frame producers
    def create_frame(self, stream, frame_data, frame_id=None):
        with self._lock:
            stream.variables["_queued_frames"] = stream.variables.get("_queued_frames", 0) + 1
        self._pending.put((stream.stream_id, stream.frame_id if frame_id is None else frame_id, frame_data))

    def _frames_generator(self, element, stream, frame_generator, frame_id, rate):
        self._enable_thread_local(stream, frame_id)
        try:
            while stream.state == StreamState.RUN:
                try:
                    event, frame_data = frame_generator(stream, frame_id)
                except Exception:
                    event, frame_data = StreamEvent.ERROR, {"diagnostic": traceback.format_exc()}
                stream.state = self._process_stream_event(element.name, stream, event, frame_data or {})
                if stream.state == StreamState.RUN and frame_data:
                    for data in ([frame_data] if isinstance(frame_data, dict) else frame_data):
                        self.create_frame(stream, data, frame_id)
                        frame_id += 1
                else:
                    frame_id += 1
                if stream.state in (StreamState.DROP_FRAME, StreamState.RUN):
                    stream.state = StreamState.RUN
                    if rate:
                        time.sleep(1.0 / rate)
                    self.thread_local.frame_id = frame_id
        finally:
            self._disable_thread_local()
            with self._lock:
                self._generators -= 1
            self._pending.put((stream.stream_id, None, None))     # generator finished

    # -- the per-frame graph walk
    def _map_in(self, element, swag):
        name = element.name
        renamed = {}
        for _, mapping in self.definition.map_in_nodes.get(name, {}).items():
            for _, to_name in mapping.items():
                renamed[to_name] = f"{name}.{to_name}"
        inputs = {}
        for spec in element.definition.input:
            key = spec["name"]
            try:
                inputs[key] = swag[renamed.get(key, key)]
            except KeyError:
                raise KeyError(f'Function parameter "{key}" not found')
        return inputs

    def _map_out(self, name, out):
        for successor, mapping in self.definition.map_out_nodes.get(name, {}).items():
            for from_name, to_name in mapping.items():
                if from_name in out:
                    out[f"{successor}.{to_name}"] = out.pop(from_name)

    def process_frame(self, stream_dict, frame_data):
        """One frame of one stream through the graph (R/main/pipeline.py:1008-1113).
        Returns (okay, swag)."""
        results = self.process_frames([(stream_dict, frame_data)])
        return results[0]

    def process_frames(self, work):
        """``work``: list of (stream_dict, frame_data), typically one frame from each of many
        streams.  Walks the graph once; at every node, elements with ``batch_across_streams``
        get all live frames in one call.  Returns [(okay, swag)] in the order of ``work``."""
        jobs = []
        for stream_dict, frame_data in work:
            stream_id = str(stream_dict.get("stream_id", DEFAULT_STREAM_ID))
            stream = self.streams.get(stream_id)
            if stream is None:
                # only the default stream is created on demand (R/main/pipeline.py:1130-1141)
                if stream_id != DEFAULT_STREAM_ID or not self.create_stream(
                        stream_id, graph_path=stream_dict.get("graph_path"), parameters=stream_dict.get("parameters")):
                    self.logger.warning(f"Process frame <{stream_id}:{stream_dict.get('frame_id')}>: stream not found")
                    jobs.append(None)
                    continue
                stream = self.streams[stream_id]
            frame_id = int(stream_dict.get("frame_id", stream.frame_id))
            stream.frame_id = frame_id
            frame = Frame()
            frame.swag.update(frame_data)
            stream.frames[frame_id] = frame
            # STOP still processes queued frames, ERROR ignores them (R/main/stream.py:48-53)
            jobs.append({"stream": stream, "frame_id": frame_id, "frame": frame,
                         "alive": stream.state in (StreamState.RUN, StreamState.STOP), "t0": time.monotonic()})
        # every stream walks the path that starts at ITS head (stream.graph_path, R/main/pipeline.py:1156);
        # frames are batched across the streams that share a path
        by_path = OrderedDict()
        for job in jobs:
            if job:
                by_path.setdefault(tuple(self.graph_path(job["stream"].graph_path)), []).append(job)
        for path, members in by_path.items():
            for node in path:
                element = self.elements[node]
                live = [j for j in members if j["alive"]]
                if not live:
                    break
                t_node = time.monotonic()
                if element.batch_across_streams and len(live) > 1:
                    self._call_batched(element, live)
                else:
                    for job in live:
                        self._call_one(element, job)
                dt = time.monotonic() - t_node
                for job in live:
                    job["frame"].metrics.setdefault("pipeline_elements", {})[f"time_{node}"] = dt
        # The graph has been walked: whatever the elements left as lazy device work is enqueued now
        # (asynchronously), so it overlaps the next frame's host work and the previous frame's read-back.
        launched = set()
        for job in jobs:
            if job and job["alive"]:
                for image in job["frame"].swag.get("images") or ():
                    batch = getattr(image, "batch", None)
                    if batch is not None and id(batch) not in launched:
                        launched.add(id(batch))
                        try:
                            batch.launch()
                        except Exception as error:   # reported when the pixels are asked for, as before
                            self.logger.debug(f"deferred launch: {error}")
        # Element times (R/main/pipeline.py:1170-1182): the host time of the call plus -- for elements whose
        # work ran lazily on the GPU -- the CUDA-event time of the kernels it caused (GpuSeconds)
        for job in jobs:
            if job and job["alive"]:
                timings = {}
                for image in job["frame"].swag.get("images") or ():
                    batch = getattr(image, "batch", None)
                    if batch is not None and getattr(batch, "timings", None):
                        timings.update(batch.timings)
                if timings:
                    elements = job["frame"].metrics.get("pipeline_elements", {})
                    for name, recorded in timings.items():
                        key = f"time_{name}"
                        if key in elements and not isinstance(elements[key], _GpuSeconds):
                            if isinstance(recorded, tuple):
                                elements[key] = _GpuSeconds(elements[key], (), fused_into=recorded[1])
                            else:
                                elements[key] = _GpuSeconds(elements[key], recorded)
        results = []
        for job in jobs:
            if job is None:
                results.append((False, {}))
                continue
            stream, frame = job["stream"], job["frame"]
            frame.metrics["time_pipeline"] = time.monotonic() - job["t0"]
            okay = job["alive"]
            if okay and stream.queue_response is not None:
                stream.queue_response.put((stream.as_dict(), frame.swag))
            stream.frames.pop(job["frame_id"], None)
            queued = stream.variables.get("_queued_frames", 0)
            if stream.state == StreamState.ERROR:
                self.destroy_stream(stream.stream_id)
            elif stream.state == StreamState.STOP and queued == 0 and not stream.frames:
                self.destroy_stream(stream.stream_id, graceful=True)      # gracefully: after queued frames
            results.append((okay, frame.swag))
        return results

    def _finish_call(self, element, job, event, out):
        stream, frame = job["stream"], job["frame"]
        state = self._process_stream_event(element.name, stream, event, out or {})
        if state == StreamState.RUN:
            out = dict(out or {})
            self._map_out(element.name, out)
            frame.swag.update(out)
        else:
            job["alive"] = False
            if state != StreamState.DROP_FRAME:
                stream.state = state

    def _call_one(self, element, job):
        stream = job["stream"]
        self._enable_thread_local(stream, job["frame_id"])
        _set_element_tag(element.name)
        try:
            inputs = self._map_in(element, job["frame"].swag)
            event, out = element.process_frame(stream, **inputs)
        except Exception:
            event, out = StreamEvent.ERROR, {"diagnostic": traceback.format_exc()}
        finally:
            self._disable_thread_local()
            _set_element_tag(None)
        self._finish_call(element, job, event, out)

    def _call_batched(self, element, live):
        calls = []
        for job in live:
            try:
                calls.append((job, self._map_in(element, job["frame"].swag)))
            except Exception:
                self._finish_call(element, job, StreamEvent.ERROR, {"diagnostic": traceback.format_exc()})
        if not calls:
            return
        _set_element_tag(element.name)
        try:
            results = element.process_frame_batch([(job["stream"], inputs) for job, inputs in calls], self)
        except Exception:
            diagnostic = {"diagnostic": traceback.format_exc()}
            results = [(StreamEvent.ERROR, diagnostic)] * len(calls)
        finally:
            _set_element_tag(None)
        for (job, _), (event, out) in zip(calls, results):
            self._finish_call(element, job, event, out)

    # -- driving frames created by elements (DataSource.create_frame / create_frames)
    def run(self, max_batch=64, idle_timeout=0.05):
        """Drain frames posted by ``create_frame(s)`` until no stream is left running and the
        queue is empty.  Frames of different streams that are ready together are processed as
        one batch (one frame per stream per graph walk keeps each stream's order)."""
        held = []                                   # frames waiting for their stream's turn
        while True:
            batch, held = held, []
            try:
                if not batch:
                    # (nobody can queue any more once the last frame generator has finished: do not wait then)
                    if self._generators == 0:
                        batch.append(self._pending.get_nowait())
                    else:
                        batch.append(self._pending.get(timeout=idle_timeout))
                while len(batch) < max_batch:
                    batch.append(self._pending.get_nowait())
            except queue.Empty:
                pass
            if not batch:
                # done when nothing is queued and nobody can queue any more: every stream has stopped,
                # or no frame generator is left running (a source that posted single frames with
                # create_frame keeps its stream in RUN; the reference ends such streams by lease expiry)
                if all(s.state != StreamState.RUN for s in self.streams.values()) or self._generators == 0:
                    for stream_id in list(self.streams):      # stopped streams with nothing queued
                        self.destroy_stream(stream_id, graceful=True)
                    return
                continue
            work, seen = [], set()
            for stream_id, frame_id, data in batch:
                if frame_id is None:                # a frame generator finished
                    continue
                if stream_id in seen:
                    held.append((stream_id, frame_id, data))
                    continue
                seen.add(stream_id)
                stream = self.streams.get(stream_id)
                if stream is not None:
                    with self._lock:
                        stream.variables["_queued_frames"] = stream.variables.get("_queued_frames", 1) - 1
                work.append(({"stream_id": stream_id, "frame_id": frame_id}, data))
            if work:
                self.process_frames(work)
