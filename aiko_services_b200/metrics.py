"""Per-element timings for ``frame.metrics`` (R/main/pipeline.py:1170-1182) when an element's work runs
lazily on the GPU.  No torch import here: events are whatever the recorder hands in."""
import threading

_current = threading.local()          # .tag: name of the element whose call is appending recipe steps


def set_element_tag(name):
    """Called by the pipeline around an element call: recipe steps appended meanwhile belong to it."""
    _current.tag = name


def current_tag():
    return getattr(_current, "tag", None)


class GpuSeconds:
    """An element's time for ``frame.metrics["pipeline_elements"]`` (R/main/pipeline.py:1170-1182):
    the host time its call took plus the GPU time of the kernels that call caused -- measured with
    CUDA events when the lazy recipe really ran, which is after the call returned.  Behaves like the
    float the reference stores (``value * 1000``, ``f"{value:.3f}"``); reading it waits for the
    kernels.  Steps that were fused into another element's kernel report that kernel under the first
    element and only their host time themselves (``fused_into`` names it)."""

    __slots__ = ("host", "events", "fused_into")

    def __init__(self, host, events=(), fused_into=None):
        self.host, self.events, self.fused_into = float(host), list(events), fused_into

    @property
    def gpu(self):
        total = 0.0
        for start, end in self.events:
            end.synchronize()
            total += start.elapsed_time(end) * 1e-3
        return total

    def __float__(self):
        return self.host + self.gpu

    def __mul__(self, other):
        return float(self) * other

    __rmul__ = __mul__

    def __add__(self, other):
        return float(self) + float(other)

    __radd__ = __add__

    def __format__(self, spec):
        return format(float(self), spec)

    def __repr__(self):
        tail = f", fused into {self.fused_into}" if self.fused_into else ""
        return f"GpuSeconds(host={self.host:.6f}, gpu={self.gpu:.6f}{tail})"
