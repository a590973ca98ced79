"""GPU: device-resident frames handed to another process through a CUDA IPC descriptor (aiko_services_b200.ipc):
the consumer process runs the hot path on the producer's frames without the pixels leaving HBM."""
import multiprocessing as mp

import numpy as np
import pytest
import torch

from golden import cases
from oracle import image as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def test_frames_cross_a_process_boundary_without_leaving_the_gpu(cuda):
    from aiko_services_b200 import ipc, ops
    from ipc_consumer import consume
    frames = np.stack([cases.frame(9100 + k, 270, 480, 3) for k in range(4)])
    producer_stream = torch.cuda.Stream(cuda)
    with torch.cuda.stream(producer_stream):
        device_frames = torch.from_numpy(frames).to(cuda, non_blocking=False)
        flipped = ops.flip(device_frames, 1, stream=producer_stream)          # producer-side work the consumer must wait for
    descriptor = ipc.export_frames(flipped, stream=producer_stream)
    assert isinstance(descriptor, bytes) and len(descriptor) < 4096            # the message, not the pixels
    ctx = mp.get_context("spawn")
    parent, child = ctx.Pipe()
    process = ctx.Process(target=consume, args=(descriptor, 160, 90, child))
    process.start()
    assert parent.poll(120), "the consumer process did not answer"
    status, payload = parent.recv()
    process.join(30)
    ipc.release(descriptor)
    assert status == "ok", payload
    for k in range(4):
        np.testing.assert_array_equal(payload[k], O.resize_linear(frames[k][:, ::-1], 160, 90)[:, :, ::-1])
