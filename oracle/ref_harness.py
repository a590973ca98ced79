"""Run the UNMODIFIED reference (read-only at /root/reference) in this container.

TEST INFRASTRUCTURE.  Used only by tests/golden/make_golden.py (to generate the committed
golden vectors) and by tests that are skipped when /root/reference is absent (the GPU box).
Nothing here is imported by the product path or smoke(); bench.py runs it only in a child
process of its CPU legs (the "as shipped" baseline: the reference's own pipeline and elements).

The reference needs six third-party modules that are not installed (paho-mqtt, avro,
avro_validator, transitions, asciimatics, pyperclip).  They are stubbed only as deep as the
reference touches them at import time; MQTT absence is something the
reference tolerates by itself (R/main/process.py:149-163, R/main/message/castaway.py).
"""
import os
import queue
import sys
import threading
import time
import types

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the read-only mount in the build container, else the git-ignored copy that travels to the GPU box
_CANDIDATES = ["/root/reference/src", os.path.join(_ROOT, "baseline", "_ref")]
REFERENCE_SRC = next((c for c in _CANDIDATES if os.path.isdir(os.path.join(c, "aiko_services"))), _CANDIDATES[0])


def available():
    return os.path.isdir(os.path.join(REFERENCE_SRC, "aiko_services"))


class _Anything:
    """Class whose every attribute is another dummy class; numeric contexts read as 0."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        return _Anything()

    def __int__(self):
        return 0

    def __index__(self):
        return 0

    def __iter__(self):
        return iter(())


class _AnyMeta(type):
    def __getattr__(cls, name):
        return 0


class _Dummy(metaclass=_AnyMeta):
    def __init__(self, *a, **k):
        pass


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Dummy


def _install_stubs():
    def mod(name, **attrs):
        m = _StubModule(name)
        m.__dict__.update(attrs)
        m.__path__ = []
        sys.modules[name] = m
        return m

    if "paho" not in sys.modules:
        class Client:
            def __init__(self, *a, **k):
                raise SystemError("paho stub: no MQTT in this container")
        mod("paho"); mod("paho.mqtt")
        mod("paho.mqtt.client", Client=Client, MQTTv311=4, CallbackAPIVersion=_Dummy)
    if "avro" not in sys.modules:
        class _Schema:
            def validate(self, datum):
                return True
        class SchemaParseException(Exception):
            pass
        avro = mod("avro")
        avro.schema = mod("avro.schema", parse=lambda s: _Schema(), SchemaParseException=SchemaParseException)
        avro.errors = mod("avro.errors", SchemaParseException=SchemaParseException)
        mod("avro_validator"); mod("avro_validator.schema", Schema=_Dummy)
    if "transitions" not in sys.modules:
        class MachineError(Exception):
            pass
        class Machine:
            def __init__(self, *a, **k):
                pass
        mod("transitions", Machine=Machine)
        mod("transitions.core", MachineError=MachineError)
    if "pyperclip" not in sys.modules:
        mod("pyperclip")
    if "asciimatics" not in sys.modules:
        mod("asciimatics")
        for sub in ("event", "exceptions", "parsers", "scene", "screen", "widgets"):
            mod(f"asciimatics.{sub}")
        mod("asciimatics.widgets.utilities", THEMES={})


_aiko = None


def load_reference():
    """import aiko_services from the reference tree (no bytecode written: tree is read-only)."""
    global _aiko
    if _aiko is None:
        if not available():
            raise RuntimeError("the reference is not present on this machine (see baseline/install_ref.sh)")
        os.environ.setdefault("AIKO_LOG_MQTT", "false")
        os.environ.setdefault("AIKO_LOG_LEVEL", "WARNING")
        sys.dont_write_bytecode = True
        _install_stubs()
        if REFERENCE_SRC not in sys.path:
            sys.path.insert(0, REFERENCE_SRC)
        import aiko_services
        _aiko = aiko_services
    return _aiko


SOURCE_ELEMENT = '''
import aiko_services as aiko
BATCHES = []          # filled by the harness: list of frame_data dicts

class SyntheticSource(aiko.PipelineElement):
    def __init__(self, context):
        context.set_protocol("synthetic_source:0")
        context.get_implementation("PipelineElement").__init__(self, context)

    def start_stream(self, stream, stream_id):
        self.create_frames(stream, self.frame_generator, rate=None)
        return aiko.StreamEvent.OKAY, {}

    def frame_generator(self, stream, frame_id):
        if frame_id < len(BATCHES):
            return aiko.StreamEvent.OKAY, BATCHES[frame_id]
        return aiko.StreamEvent.STOP, {"diagnostic": "All frames generated"}

    def process_frame(self, stream, **inputs):
        return aiko.StreamEvent.OKAY, inputs
'''


def run_reference_pipeline(graph_elements, batches, parameters=None, workdir="/tmp/aiko_ref_harness",
                           timeout=120.0, timestamps=None):
    """Drive the reference's own PipelineImpl over ``batches`` (list of frame_data dicts).

    graph_elements: list of (element_name, module, inputs, outputs, element_parameters), in
    graph order, placed after the synthetic source.  Returns the list of frame_data_out dicts
    the pipeline put on its queue_response, in frame order.  ``timestamps`` (a list) receives the
    perf_counter time at which each response arrived (the as-shipped CPU baseline of bench.py).
    """
    import json
    aiko = load_reference()
    from aiko_services.main.pipeline import PipelineImpl
    os.makedirs(workdir, exist_ok=True)
    src_path = os.path.join(workdir, "synthetic_source.py")
    with open(src_path, "w") as f:
        f.write(SOURCE_ELEMENT)
    keys = sorted({k for b in batches for k in b})
    io = [{"name": k, "type": "any"} for k in keys]
    names = ["SyntheticSource"] + [g[0] for g in graph_elements]
    elements = [{"name": "SyntheticSource", "input": io, "output": io,
                 "deploy": {"local": {"module": src_path}}}]
    for name, module, inputs, outputs, params in graph_elements:
        e = {"name": name, "input": [{"name": k, "type": "any"} for k in inputs],
             "output": [{"name": k, "type": "any"} for k in outputs],
             "deploy": {"local": {"module": module}}}
        if params:
            e["parameters"] = params
        elements.append(e)
    definition = {"version": 0, "name": "p_harness", "runtime": "python",
                  "graph": ["(" + " ".join(names) + ")"], "parameters": parameters or {},
                  "elements": elements}
    json_path = os.path.join(workdir, "pipeline.json")
    with open(json_path, "w") as f:
        json.dump(definition, f)

    from aiko_services.main.utilities import importer
    importer.MODULES_LOADED.pop(src_path, None)
    module = importer.load_module(src_path)
    module.BATCHES[:] = batches

    responses = queue.Queue()
    results = []
    done = threading.Event()

    def watcher():
        while len(results) < len(batches):
            try:
                item = responses.get(timeout=timeout)
            except queue.Empty:
                break
            results.append(item)
            if timestamps is not None:
                timestamps.append(time.perf_counter())
        done.set()
        aiko.process.terminate()

    defn = PipelineImpl.parse_pipeline_definition(json_path)
    pipeline = PipelineImpl.create_pipeline(json_path, defn, None, None, "1", [], 0, None, 60,
                                            queue_response=responses)
    threading.Thread(target=watcher, daemon=True).start()
    pipeline.run(mqtt_connection_required=False)
    done.wait(timeout)
    out = []
    for item in results:
        stream_info, frame_data = item
        out.append((stream_info, frame_data))
    out.sort(key=lambda t: int(t[0]["frame_id"]) if isinstance(t[0], dict) else 0)
    return [fd for _, fd in out]
