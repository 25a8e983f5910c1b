"""CUDA-graph front end of the render step.

One inversion step launches ~40 of this library's kernels plus ~60 small ATen ones; only the four
field launches are long (2.8 / 6.9 ms), the rest run for 5-500 us each -- shorter than it takes
Python to launch them -- so an eager step leaves the GPU idle for about a millisecond.  The shapes
of the inversion loop are static for all `inv_steps` iterations (run.py:2256-2317), so the whole
step -- render forward, loss, backward -- is captured once and replayed:

    step = GraphedStep(fn, example_inputs)      # fn(*tensors) -> tensor or tuple of tensors
    outs = step(*new_inputs)                    # values are copied into the captured input buffers
    outs = step()                               # or: update step.inputs[i] in place, then replay

`fn` must be capturable: no host synchronisation (render with `check_hits="defer"`; the deferred
"no ray hits the box" check of lib/nerf_utils.py:260 is resolved by `step.check_hits()`), static
shapes, all work on the current stream.  torch.rand inside `fn` is graph-safe (the Philox offset
advances per replay), so every replay draws fresh uniforms exactly like an eager step.
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch

from . import ops


class GraphedStep:
    def __init__(self, fn: Callable, example_inputs: Sequence[torch.Tensor], warmup: int = 3):
        assert len(example_inputs) > 0 and all(t.is_cuda for t in example_inputs), \
            "GraphedStep: CUDA tensors only (there is no CPU path)"
        self.device = example_inputs[0].device
        self.inputs = [t.detach().clone().requires_grad_(t.requires_grad) for t in example_inputs]
        self._pending = []
        # warm-up on a side stream (allocator + lazy initialisation must not happen under capture)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                fn(*self.inputs)
        torch.cuda.current_stream(self.device).wait_stream(side)
        ops.raise_if_no_hits()                      # the warm-up's deferred checks
        before = ops.count_launches()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            out = fn(*self.inputs)
        self.launches_per_replay = ops.count_launches() - before
        self._pending = ops.take_pending_hits()     # device counters rewritten by every replay
        self._single = torch.is_tensor(out)
        self.outputs = (out,) if self._single else tuple(out)

    def __call__(self, *inputs):
        if inputs:
            assert len(inputs) == len(self.inputs)
            with torch.no_grad():
                for dst, src in zip(self.inputs, inputs):
                    if src is not dst:
                        dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.outputs[0] if self._single else self.outputs

    def check_hits(self):
        """Raise like the reference (nerf_utils.py:260) if the last replay had no ray hitting the box.
        One small device->host read; call it at a point where the caller synchronises anyway."""
        for t in self._pending:
            if int(t.min().item()) == 0:
                raise RuntimeError(ops._NO_HIT_MSG)
