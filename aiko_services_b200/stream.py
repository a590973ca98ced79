"""Stream / Frame API types -- same names, fields and constants as the reference's
R/main/stream.py:33-98, so elements written against either package read the same."""
from dataclasses import dataclass, field
from typing import Any, Dict

__all__ = ["DEFAULT_STREAM_ID", "FIRST_FRAME_ID", "Frame", "Stream", "StreamEvent", "StreamEventName",
           "StreamState", "StreamStateName"]

DEFAULT_STREAM_ID = "*"
FIRST_FRAME_ID = 0


class StreamEvent:            # R/main/stream.py:33-38
    ERROR = -2
    STOP = -1
    OKAY = 0
    DROP_FRAME = 1
    USER = 1024


StreamEventName = {StreamEvent.DROP_FRAME: "DropFrame", StreamEvent.ERROR: "Error", StreamEvent.OKAY: "Okay",
                   StreamEvent.STOP: "Stop", StreamEvent.USER: "User"}


class StreamState:            # R/main/stream.py:48-53
    ERROR = -2
    STOP = -1
    RUN = 0
    DROP_FRAME = 1
    USER = 1024


StreamStateName = {StreamState.DROP_FRAME: "DropFrame", StreamState.ERROR: "Error", StreamState.STOP: "Stop",
                   StreamState.RUN: "Run", StreamState.USER: "User"}


@dataclass
class Frame:                  # R/main/stream.py:63-69
    metrics: Dict[str, Any] = field(default_factory=dict)
    paused_pe_name: str = None
    swag: Dict[str, Any] = field(default_factory=dict)


@dataclass
class Stream:                 # R/main/stream.py:71-98
    stream_id: str = DEFAULT_STREAM_ID
    frame_id: int = FIRST_FRAME_ID
    graph_path: str = None
    frames: Dict[int, Frame] = field(default_factory=dict)
    parameters: Dict[str, Any] = field(default_factory=dict)
    queue_response: Any = None
    state: int = StreamState.RUN
    topic_response: str = None
    variables: Dict[str, Any] = field(default_factory=dict)

    def as_dict(self):
        return {"stream_id": self.stream_id, "frame_id": self.frame_id}

    def update(self, stream_dict):
        if not isinstance(stream_dict, dict):
            return False
        self.stream_id = str(stream_dict.get("stream_id", self.stream_id))
        self.frame_id = int(stream_dict.get("frame_id", self.frame_id))
        self.graph_path = stream_dict.get("graph_path", self.graph_path)
        self.parameters = stream_dict.get("parameters", self.parameters)
        self.state = int(stream_dict.get("state", StreamState.RUN))
        return True
