"""CUDA-graph capture of a whole render step (forward chain, loss, backward chain: ~11 kernels of libtsplat.so).

Eagerly a C2 step costs the host 0.3-0.45 ms of Python + launch work against 0.78 ms of device time, which is fine
for one process per box but not for 8 ranks on a 16-vCPU host: measured 1.27 ms/step/rank at 8 GPUs, host-bound.
A captured step is ONE graph launch.  The library itself is capture-safe (it only enqueues kernels on the stream it
is given); under capture ``rasterize`` skips everything that would touch the host (status copy, event polling,
capacity growth), so one eager call of the same problem shape must come first and the caller checks
``read_status_device`` after replays.  PyTorch is plumbing here: ``torch.cuda.CUDAGraph`` owns the capture, the
private memory pool and the replay.
"""
from __future__ import annotations

from typing import Callable

import torch


class GraphedStep:
    """``fn()`` reads tensors at fixed addresses (its inputs live outside, e.g. resident scene tensors or the device
    slots of a HostPipeline) and returns tensors; after construction ``replay()`` re-runs the captured kernels and the
    returned ``outputs`` are refreshed in place."""

    def __init__(self, fn: Callable[[], object], warmup: int = 2, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):  # eager warm-up on a side stream: calibrates capacities, builds autograd state
            for _ in range(max(1, warmup)):
                out = fn()
            del out
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.outputs = fn()

    def replay(self):
        self.graph.replay()
        return self.outputs
