"""CUDA-graph replay of a hot-path step.

A step of the path is 10-25 short launches (assignment, list, stream, serve, scale; score, the NMS chain,
gather) next to two long streaming kernels: at small per-GPU batches the host, not the GPU, sets the pace
(C3: 0.17 ms of kernels in a 0.27 ms FocalLoss forward+backward).  Everything the library launches is
capturable - no allocation, no synchronisation, the assignment's generation number and the peer exchange's
call number are device-side under capture (K7LossParams.consume_best, K7PeerReduce.seq_counter) - so a step
over STATIC tensors (the usual whole-step capture of a training loop) can be recorded once and replayed.
"""
from __future__ import annotations

import torch

from . import _lib as L


class GraphedStep:
    """``GraphedStep(fn)`` runs ``fn()`` a few times on a private stream (workspaces are per stream: they are
    created here, outside the graph), captures one more call in a CUDA graph - ``fn`` may fork work onto other
    streams as long as it joins them again - and replays it on every ``__call__``.  ``fn`` must read and write
    the same tensors every time (copy new inputs into them before replaying).  ``launches`` is the number of
    library kernels one replay launches (``k7_launch_count`` over the capture)."""

    def __init__(self, fn, warmup: int = 3, device=None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.stream = torch.cuda.Stream(device=dev)
        cur = torch.cuda.current_stream(dev)
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            for _ in range(max(1, warmup)):
                fn()
        cur.wait_stream(self.stream)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        n0 = L.lib().k7_launch_count()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self.result = fn()
        self.launches = int(L.lib().k7_launch_count() - n0)

    def __call__(self):
        self.graph.replay()
        return self.result
