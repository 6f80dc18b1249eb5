"""Whole-forward CUDA graphs.

RAFT's refinement loop (core/raft.py:122-139) launches a few hundred small kernels per frame pair; at batch 1 the
GPU waits for the host between them.  Every kernel of this library is capturable (no allocation, no implicit
synchronisation, stream passed in), so a complete no-grad forward — encoders, volume build, the 12 lookups, update
blocks, upsampling — can be recorded once and replayed as ONE graph launch:

    step = raftcorr_b200.graphed(lambda a, b: model(a, b, iters=12, test_mode=True), (image1, image2))
    flow_low, flow_up = step(image1, image2)        # copies the inputs in, replays, returns the graph's outputs

The callable must be shape-static and free of host-side data dependence (true for raft.py in test_mode).  Outputs
are the graph's own buffers: they are overwritten by the next call (``clone`` them to keep them).
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch


class GraphedForward:
    def __init__(self, fn: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 2):
        if not example_inputs or not all(t.is_cuda for t in example_inputs):
            raise RuntimeError("graphed() needs CUDA example inputs")
        self._device = example_inputs[0].device
        self._inputs = [t.detach().clone() for t in example_inputs]
        self._graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(self._device)
        side.wait_stream(torch.cuda.current_stream(self._device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(warmup, 1)):          # one-time set-up (kernel attributes, tensor-map caches, cuDNN plans)
                fn(*self._inputs)                    # must happen outside the capture
            side.synchronize()
            with torch.cuda.graph(self._graph, stream=side):
                self._outputs = fn(*self._inputs)
        torch.cuda.current_stream(self._device).wait_stream(side)

    def __call__(self, *inputs: torch.Tensor):
        if len(inputs) != len(self._inputs):
            raise RuntimeError(f"expected {len(self._inputs)} inputs, got {len(inputs)}")
        for dst, src in zip(self._inputs, inputs):
            if dst.shape != src.shape or dst.dtype != src.dtype:
                raise RuntimeError(f"graphed forward was captured for {tuple(dst.shape)} {dst.dtype}, "
                                   f"got {tuple(src.shape)} {src.dtype}")
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self._graph.replay()
        return self._outputs


def graphed(fn: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 2) -> GraphedForward:
    """Capture ``fn(*example_inputs)`` (run under ``torch.no_grad``) as one CUDA graph; returns a callable."""
    return GraphedForward(fn, example_inputs, warmup)
