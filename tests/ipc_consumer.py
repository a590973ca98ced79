"""The receiving process of tests/test_gpu_ipc.py: maps the exported frames and runs the hot path on them."""


def consume(descriptor, width, height, connection):
    import numpy as np
    import torch

    from aiko_services_b200 import ipc, ops
    try:
        batch = ipc.import_frames(descriptor)
        with torch.cuda.stream(batch.stream):
            out = ops.resize_linear(batch.tensor, width, height, swap_rb=True, stream=batch.stream)
            host = out.cpu()
        batch.stream.synchronize()
        connection.send(("ok", np.ascontiguousarray(host.numpy())))
    except Exception as error:       # the parent fails the test with the message
        connection.send(("error", repr(error)))
    finally:
        connection.close()
