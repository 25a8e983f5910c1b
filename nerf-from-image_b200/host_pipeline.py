"""Host-buffer front end of the render step: images are independent, so a batch
that lives in pinned host memory is cut into chunks of images and the three legs of a chunk --
host->device copy of its inputs, render forward + backward, device->host copy of its outputs and
gradients -- run on three CUDA streams, chained by events.  PCIe is full duplex: while chunk i
computes, chunk i+1 is arriving and chunk i-1 is leaving.

The per-chunk evaluation of the miss-ray near/far fill (lib/nerf_utils.py:260-261) is the same
liberty nn.DataParallel takes in the reference (one evaluation per replica chunk, run.py:636-644);
it never reaches rgb / depth / mask.
"""
from typing import Callable, Sequence

import torch


class HostStepPipeline:
    """step_fn(*device_inputs) -> sequence of device tensors; outputs with a leading batch dimension
    are written to the matching rows of the host output, 0-dim outputs (losses) are summed."""

    def __init__(self, step_fn: Callable, batch: int, n_chunks: int, device: torch.device):
        if batch % n_chunks != 0:
            raise ValueError("HostStepPipeline: batch %d is not a multiple of %d chunks" % (batch, n_chunks))
        self.step_fn = step_fn
        self.batch = batch
        self.n_chunks = n_chunks
        self.per = batch // n_chunks
        self.device = device
        self.s_in = torch.cuda.Stream(device)
        self.s_out = torch.cuda.Stream(device)

    def _slice(self, t: torch.Tensor, c: int):
        if t.dim() > 0 and t.shape[0] == self.batch:
            return t[c * self.per:(c + 1) * self.per]
        return t

    def run(self, host_in: Sequence[torch.Tensor], host_out: Sequence[torch.Tensor]):
        """host_in / host_out: pinned CPU tensors.  Returns after every host output is complete."""
        comp = torch.cuda.current_stream(self.device)
        n = self.n_chunks
        ev_in = [torch.cuda.Event() for _ in range(n)]
        ev_done = [torch.cuda.Event() for _ in range(n)]
        keep = []                                   # device tensors stay referenced until the final sync
        chunks = []
        self.s_in.wait_stream(comp)
        with torch.cuda.stream(self.s_in):
            for c in range(n):
                dev_in = [self._slice(t, c).to(self.device, non_blocking=True) for t in host_in]
                ev_in[c].record(self.s_in)
                chunks.append(dev_in)
        keep.append(chunks)
        scalars = [None] * len(host_out)
        for c in range(n):
            comp.wait_event(ev_in[c])
            outs = self.step_fn(*chunks[c])
            ev_done[c].record(comp)
            keep.append(outs)
            self.s_out.wait_event(ev_done[c])
            with torch.cuda.stream(self.s_out):
                for k, (dst, src) in enumerate(zip(host_out, outs)):
                    src = src.detach()
                    if dst.dim() == 0:
                        scalars[k] = src if scalars[k] is None else scalars[k] + src
                        if c == n - 1:
                            dst.copy_(scalars[k], non_blocking=True)
                    else:
                        self._slice(dst, c).copy_(src, non_blocking=True)
        self.s_out.synchronize()
        comp.wait_stream(self.s_out)
        del keep
        from . import ops
        ops.raise_if_no_hits()          # deferred "no ray hits the box" checks of the chunks
