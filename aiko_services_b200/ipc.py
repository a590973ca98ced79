"""Device-resident frames handed to ANOTHER PROCESS on the same box without leaving HBM.

The reference sends a frame to an element that is deployed remotely by serialising its data into the actor
message that invokes ``process_frame`` in the other process (R/main/pipeline.py:1080-1090, 1285-1319;
consumer side R/examples/pipeline/elements.py:298-324): every pixel crosses the message transport.  For two
processes that share a GPU (or NVLink-connected GPUs of one box) a 1080p batch is 199 MB of pixels but its CUDA
IPC handle is 64 bytes: ``export_frames`` turns a ``DeviceBatch`` / CUDA tensor into a small bytes descriptor
that fits any message, ``import_frames`` maps the same memory in the receiving process and returns a
``DeviceBatch`` whose frames downstream elements use like any other.  Ordering: the descriptor carries an
interprocess CUDA event recorded after the producer's work; the consumer's stream waits on it.

Only the transport is here -- the control plane that creates remote element proxies and carries the message
(actors over MQTT) is outside this package's scope.  The producer must keep the exported frames alive and
unmodified until the consumer is done with them (the handle does not own the memory), e.g. until the reply to
the message arrives; ``release`` drops the producer's keep-alive reference.
"""
import pickle

import torch

from .device_frames import DeviceBatch

_exports = {}          # token -> (tensor, event): keeps exported memory alive in the producer
_next_token = [0]


def export_frames(frames, stream=None):
    """``frames``: a DeviceBatch (materialised here, on its own stream) or a contiguous CUDA tensor.
    -> bytes: a descriptor for ``import_frames`` in another process on the same box."""
    from torch.multiprocessing.reductions import reduce_tensor
    if isinstance(frames, DeviceBatch):
        stream = frames.stream if frames.stream is not None else stream
        tensor = frames.tensor
    else:
        tensor = frames
    if not (isinstance(tensor, torch.Tensor) and tensor.is_cuda and tensor.is_contiguous()):
        raise RuntimeError("export_frames: a contiguous CUDA tensor or a DeviceBatch is needed")
    s = stream if stream is not None else torch.cuda.current_stream(tensor.device)
    event = torch.cuda.Event(interprocess=True)
    event.record(s)                                    # everything the producer queued for these frames
    _rebuild, args = reduce_tensor(tensor)             # CUDA IPC handle of the allocation + offsets, shape, strides
    token = _next_token[0]
    _next_token[0] += 1
    _exports[token] = (tensor, event)
    return pickle.dumps({"token": token, "tensor": args, "event": event.ipc_handle(), "device": tensor.device.index})


def import_frames(descriptor, stream=None):
    """-> DeviceBatch over the producer's memory (no copy).  The work of ``stream`` (default: a new CUDA stream
    of the frames' device) is ordered after the producer's."""
    from torch.multiprocessing.reductions import rebuild_cuda_tensor
    d = pickle.loads(descriptor)
    device = torch.device("cuda", d["device"])
    with torch.cuda.device(device):
        tensor = rebuild_cuda_tensor(*d["tensor"])
        s = stream if stream is not None else torch.cuda.Stream(device)
        event = torch.cuda.Event.from_ipc_handle(device, d["event"])
        s.wait_event(event)
    batch = DeviceBatch(tensor, s)
    batch.ipc_token = d["token"]
    return batch


def release(descriptor_or_token):
    """Producer side: the consumer is done with an export (its reply arrived); drop the keep-alive reference."""
    token = descriptor_or_token if isinstance(descriptor_or_token, int) else pickle.loads(descriptor_or_token)["token"]
    _exports.pop(token, None)
