"""CPU: host-side logic of the runtime -- definition parsing, graph order, parameter
resolution, stream events, cross-stream batching -- with plain Python elements (no GPU)."""
import json
import os
import queue

import pytest

from aiko_services_b200.pipeline import Pipeline, PipelineElement, execution_order, parse_graph, parse_pipeline_definition
from aiko_services_b200.stream import Frame, Stream, StreamEvent, StreamState

ELEMENTS = '''
from aiko_services_b200.pipeline import PipelineElement
from aiko_services_b200.stream import StreamEvent

class PE_Add(PipelineElement):
    def process_frame(self, stream, i):
        constant, _ = self.get_parameter("constant", default=1)
        return StreamEvent.OKAY, {"i": int(i) + int(constant)}

class PE_Rename(PipelineElement):
    def process_frame(self, stream, x):
        return StreamEvent.OKAY, {"y": x * 10}

class PE_Take(PipelineElement):
    def process_frame(self, stream, z):
        return StreamEvent.OKAY, {"z": z + 1}

class PE_Event(PipelineElement):
    def process_frame(self, stream, i):
        if i == 13: raise ValueError("unlucky")
        if i == 7:  return StreamEvent.DROP_FRAME, {}
        if i == 99: return StreamEvent.STOP, {"diagnostic": "done"}
        if i == -1: return StreamEvent.ERROR, {"diagnostic": "bad"}
        return StreamEvent.OKAY, {"i": i}

class PE_Batch(PipelineElement):
    batch_across_streams = True
    calls = []
    def process_frame(self, stream, i):
        PE_Batch.calls.append(("one", stream.stream_id))
        return StreamEvent.OKAY, {"i": i * 2}
    def process_frame_batch(self, calls, pipeline):
        PE_Batch.calls.append(("batch", [s.stream_id for s, _ in calls]))
        return [(StreamEvent.OKAY, {"i": inputs["i"] * 2}) for _, inputs in calls]

class PE_Source(PipelineElement):
    def start_stream(self, stream, stream_id):
        self.create_frames(stream, self.frame_generator)
        return StreamEvent.OKAY, {}
    def frame_generator(self, stream, frame_id):
        limit, _ = self.get_parameter("limit", default=3)
        if frame_id < int(limit):
            return StreamEvent.OKAY, {"i": frame_id}
        return StreamEvent.STOP, {"diagnostic": "All frames generated"}
    def process_frame(self, stream, i):
        return StreamEvent.OKAY, {"i": i}
'''


@pytest.fixture()
def module_path(tmp_path):
    path = tmp_path / "elements.py"
    path.write_text(ELEMENTS)
    return str(path)


def definition(module_path, graph, names, parameters=None, element_parameters=None, io=None):
    io = io or {}
    elements = []
    for name in names:
        class_name = name.split("_x")[0]
        i, o = io.get(name, (["i"], ["i"]))
        e = {"name": name, "input": [{"name": k, "type": "int"} for k in i],
             "output": [{"name": k, "type": "int"} for k in o],
             "deploy": {"local": {"module": module_path, "class_name": class_name}}}
        if element_parameters and name in element_parameters:
            e["parameters"] = element_parameters[name]
        elements.append(e)
    return {"version": 0, "name": "p_test", "runtime": "python", "graph": graph,
            "parameters": parameters or {}, "elements": elements}


def test_graph_orders():
    h, s, r = parse_graph(["(A B C)"])
    assert execution_order(h, s) == ["A", "B", "C"]
    h, s, r = parse_graph(["(A (B D) C)"])
    assert execution_order(h, s) == ["A", "B", "D", "C"]
    h, s, r = parse_graph(["(A (B D) (C D))"])
    assert execution_order(h, s) == ["A", "B", "C", "D"]      # a node reached again moves later
    h, s, r = parse_graph(["(A B (x: y) C)"])
    assert r == [("A", "B", {"x": "y"})]


def test_definition_validation(module_path, tmp_path):
    d = definition(module_path, ["(PE_Add)"], ["PE_Add"])
    path = tmp_path / "p.json"
    path.write_text(json.dumps(d))
    assert parse_pipeline_definition(str(path)).name == "p_test"
    for broken in ({**d, "version": 1}, {**d, "runtime": "go"}, {k: v for k, v in d.items() if k != "graph"}):
        with pytest.raises(SystemExit):
            parse_pipeline_definition(broken)
    with pytest.raises(SystemExit):
        Pipeline(definition(module_path, ["(PE_Add PE_Missing)"], ["PE_Add"]))


def test_chain_and_parameter_resolution(module_path):
    d = definition(module_path, ["(PE_Add PE_Add_x2)"], ["PE_Add", "PE_Add_x2"], parameters={"constant": 5},
                   element_parameters={"PE_Add_x2": {"constant": 100}})
    p = Pipeline(d)
    q = queue.Queue()
    assert p.create_stream("s1", queue_response=q)
    okay, swag = p.process_frame({"stream_id": "s1", "frame_id": 0}, {"i": 1})
    assert okay and swag["i"] == 1 + 5 + 100          # pipeline parameter, then element parameter
    assert q.get_nowait()[1]["i"] == 106
    # stream parameters override: "Element.name" first, then bare name (pipeline.py:422-456)
    assert p.create_stream("s2", parameters={"PE_Add_x2.constant": 1, "constant": 2})
    okay, swag = p.process_frame({"stream_id": "s2"}, {"i": 0})
    assert swag["i"] == 0 + 2 + 1
    assert p.destroy_stream("s1") and not p.destroy_stream("s1")


def test_rename_mapping(module_path):
    d = definition(module_path, ["(PE_Rename PE_Take (y: z))"], ["PE_Rename", "PE_Take"],
                   io={"PE_Rename": (["x"], ["y"]), "PE_Take": (["z"], ["z"])})
    p = Pipeline(d)
    okay, swag = p.process_frame({"stream_id": "*"}, {"x": 4})      # the default stream is created on demand
    assert okay and swag["z"] == 41
    okay, _ = p.process_frame({"stream_id": "1"}, {"x": 4})         # any other must be created first
    assert not okay


def test_stream_events(module_path):
    p = Pipeline(definition(module_path, ["(PE_Event PE_Add)"], ["PE_Event", "PE_Add"]))
    for sid in "abcd":
        assert p.create_stream(sid)
    okay, swag = p.process_frame({"stream_id": "a"}, {"i": 1})
    assert okay and swag["i"] == 2
    okay, _ = p.process_frame({"stream_id": "a", "frame_id": 1}, {"i": 7})          # DROP_FRAME: stream lives on
    assert not okay and "a" in p.streams and p.streams["a"].state == StreamState.RUN
    okay, _ = p.process_frame({"stream_id": "a", "frame_id": 2}, {"i": 13})         # exception -> ERROR -> destroyed
    assert not okay and "a" not in p.streams
    okay, _ = p.process_frame({"stream_id": "b"}, {"i": -1})
    assert not okay and "b" not in p.streams
    okay, _ = p.process_frame({"stream_id": "c"}, {"i": 99})                        # STOP: graceful destroy
    assert "c" not in p.streams
    okay, _ = p.process_frame({"stream_id": "d"}, {"missing": 1})                   # input not in swag -> ERROR
    assert not okay


def test_cross_stream_batching(module_path):
    p = Pipeline(definition(module_path, ["(PE_Add PE_Batch)"], ["PE_Add", "PE_Batch"]))
    import importlib  # the module was loaded from a path: fetch the class through an element
    cls = type(p.elements["PE_Batch"])
    cls.calls.clear()
    for k in range(8):
        p.create_stream(f"s{k:02d}")
    work = [({"stream_id": f"s{k:02d}", "frame_id": 0}, {"i": k}) for k in range(8)]
    results = p.process_frames(work)
    assert [swag["i"] for _, swag in results] == [(k + 1) * 2 for k in range(8)]
    assert cls.calls == [("batch", [f"s{k:02d}" for k in range(8)])]               # ONE call for 8 streams
    cls.calls.clear()
    p.process_frame({"stream_id": "s00", "frame_id": 1}, {"i": 1})
    assert cls.calls == [("one", "s00")]


def test_generator_driven_run(module_path):
    p = Pipeline(definition(module_path, ["(PE_Source PE_Add)"], ["PE_Source", "PE_Add"],
                            element_parameters={"PE_Source": {"limit": 4}}))
    q = queue.Queue()
    for sid in ("x", "y"):
        p.create_stream(sid, queue_response=q)
    p.run()
    got = {}
    while not q.empty():
        info, swag = q.get()
        got.setdefault(info["stream_id"], []).append(swag["i"])
    assert got == {"x": [1, 2, 3, 4], "y": [1, 2, 3, 4]}


def test_stream_types_match_reference_constants():
    assert (StreamEvent.ERROR, StreamEvent.STOP, StreamEvent.OKAY, StreamEvent.DROP_FRAME, StreamEvent.USER) == \
        (-2, -1, 0, 1, 1024)
    s = Stream()
    assert s.stream_id == "*" and s.frame_id == 0 and s.as_dict() == {"stream_id": "*", "frame_id": 0}
    assert s.update({"stream_id": 5, "frame_id": "3"}) and s.stream_id == "5" and s.frame_id == 3
    assert not s.update("nope") and Frame().swag == {}


def test_reference_modules_resolve_to_gpu_elements():
    """Existing JSON names aiko_services.elements.media.image_io: this runtime serves it."""
    from aiko_services_b200.pipeline import _load_module
    module = _load_module("aiko_services.elements.media.image_io")
    assert module.__name__ == "aiko_services_b200.elements.media.image_io"
    for name in ("ImageOutput", "ImageOverlay", "ImageReadFile", "ImageResize", "ImageWriteFile"):
        assert issubclass(getattr(module, name), PipelineElement)


def test_streams_walk_the_path_of_their_own_head(tmp_path):
    """Two heads: a stream created with graph_path=B gets start_stream on B's elements AND its frames
    run through B's elements (R/main/pipeline.py:1156 walks get_path(stream.graph_path) per frame)."""
    import textwrap
    from aiko_services_b200.pipeline import Pipeline
    mod = tmp_path / "two_heads.py"
    mod.write_text(textwrap.dedent("""
        from aiko_services_b200.pipeline import PipelineElement
        from aiko_services_b200.stream import StreamEvent

        class Tag(PipelineElement):
            def start_stream(self, stream, stream_id):
                stream.variables.setdefault("started", []).append(self.name)
                return StreamEvent.OKAY, None
            def process_frame(self, stream, trail):
                return StreamEvent.OKAY, {"trail": trail + [self.name]}
    """))
    io = [{"name": "trail", "type": "list"}]
    element = lambda name: {"name": name, "input": io, "output": io,
                            "deploy": {"local": {"module": str(mod), "class_name": "Tag"}}}
    p = Pipeline({"version": 0, "name": "p_heads", "runtime": "python", "graph": ["(A A2)", "(B B2)"],
                  "elements": [element(n) for n in ("A", "A2", "B", "B2")]})
    assert p.create_stream("sa") and p.create_stream("sb", graph_path="B")
    assert p.streams["sa"].variables["started"] == ["A", "A2"]
    assert p.streams["sb"].variables["started"] == ["B", "B2"]
    results = p.process_frames([({"stream_id": "sa", "frame_id": 0}, {"trail": []}),
                                ({"stream_id": "sb", "frame_id": 0}, {"trail": []})])
    assert [r[1]["trail"] for r in results] == [["A", "A2"], ["B", "B2"]]


def test_locators_items_and_batches(tmp_path):
    """common_io: URL parsing, "{}" patterns, ordering, batches, diagnostics."""
    from aiko_services_b200.elements.media.common_io import (Batcher, Locator, LocatorError, contains_all,
                                                              file_glob_difference)
    for k in (2, 0, 1):
        (tmp_path / f"in_{k:02d}.jpeg").write_bytes(b"x")
    (tmp_path / "other.txt").write_bytes(b"y")
    loc = Locator.parse(f"file://{tmp_path}/in_{{}}.jpeg")
    assert loc.location == str(tmp_path) and loc.pattern == "in_*.jpeg"
    items = loc.expand()
    assert [i.path.name for i in items] == ["in_00.jpeg", "in_01.jpeg", "in_02.jpeg"]
    assert [i.file_id for i in items] == ["00", "01", "02"]
    assert [i.path.name for i in Locator.parse(str(tmp_path / "other.txt")).expand()] == ["other.txt"]
    assert len(Locator.parse(str(tmp_path)).expand()) == 4 and Locator.parse(str(tmp_path)).expand()[0].file_id is None
    both = Locator.parse_list(f"(file://{tmp_path}/in_{{}}.jpeg {tmp_path}/other.txt)")
    assert len(both) == 2
    with pytest.raises(LocatorError, match="scheme"):
        Locator.parse("http://example/x")
    with pytest.raises(LocatorError, match="does not exist"):
        Locator.parse(str(tmp_path / "missing")).expand()
    b = Batcher(items)
    assert [len(b.take(2)), len(b.take(2)), len(b.take(2))] == [2, 1, 0]
    assert contains_all("a{}b", "{}") and not contains_all("ab", "{}")
    assert file_glob_difference("in_*.jpeg", "in_07.jpeg") == "07" and file_glob_difference("in_*.jpeg", "x.png") is None
