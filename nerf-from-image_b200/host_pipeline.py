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

    def __init__(self, step_fn: Callable, batch: int, n_chunks, device: torch.device):
        """n_chunks: number of equal chunks, or a sequence of chunk sizes (images) that sums to batch --
        small first and last chunks shorten the copy-in / copy-out that nothing overlaps."""
        if isinstance(n_chunks, int):
            if batch % n_chunks != 0:
                raise ValueError("HostStepPipeline: batch %d is not a multiple of %d chunks" % (batch, n_chunks))
            sizes = [batch // n_chunks] * n_chunks
        else:
            sizes = [int(v) for v in n_chunks]
            if sum(sizes) != batch or min(sizes) <= 0:
                raise ValueError("HostStepPipeline: chunk sizes %s do not partition a batch of %d" % (sizes, batch))
        self.step_fn = step_fn
        self.batch = batch
        self.sizes = sizes
        self.n_chunks = len(sizes)
        self.starts = [sum(sizes[:i]) for i in range(len(sizes))]
        self.device = device
        self.s_in = torch.cuda.Stream(device)
        self.s_out = torch.cuda.Stream(device)

    def _slice(self, t: torch.Tensor, c: int):
        if t.dim() > 0 and t.shape[0] == self.batch:
            return t[self.starts[c]:self.starts[c] + self.sizes[c]]
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


class GraphedHostStepPipeline(HostStepPipeline):
    """Same three-stream pipeline with the chunk step captured as CUDA graphs (nfi_b200.graph).

    A chunk of 1-2 images keeps the GPU busy for 1-3 ms, about as long as Python needs to launch its
    ~100 kernels, so the eager pipeline is launch-bound.  Here `depth` GraphedStep instances of
    `step_fn` are captured once on chunk-shaped buffers and used round-robin: the host->device copy
    of chunk c lands directly in the input buffers of graph c % depth (free once chunk c - depth has
    been replayed), the replay writes that graph's output buffers (free once chunk c - depth has
    left for the host).  `step_fn` must be capturable (render with check_hits="defer").
    """

    def __init__(self, step_fn: Callable, batch: int, n_chunks, device: torch.device,
                 example_host_in: Sequence[torch.Tensor], depth: int = 2):
        super().__init__(step_fn, batch, n_chunks, device)
        from .graph import GraphedStep
        self.depth = max(1, depth)
        self.pools = {}                             # chunk size -> graph instances used round-robin
        self.inst = []                              # chunk -> its instance
        used = {}
        for c, size in enumerate(self.sizes):
            pool = self.pools.setdefault(size, [])
            k = used.get(size, 0)
            used[size] = k + 1
            if len(pool) < self.depth and k >= len(pool):
                pool.append(GraphedStep(step_fn, [self._slice(t, c).to(device) for t in example_host_in]))
            self.inst.append(pool[k % len(pool)])
        self.launches_per_chunk = {size: pool[0].launches_per_replay for size, pool in self.pools.items()}
        pend = [t for pool in self.pools.values() for g in pool for t in g._pending]
        self._hit_min = torch.ones((), dtype=pend[0].dtype, device=device) if pend else None

    def run(self, host_in: Sequence[torch.Tensor], host_out: Sequence[torch.Tensor]):
        comp = torch.cuda.current_stream(self.device)
        n, d = self.n_chunks, self.depth
        ev_in = [torch.cuda.Event() for _ in range(n)]
        last_done, last_out = {}, {}                # instance -> events of the chunk that used it last
        self.s_in.wait_stream(comp)
        if self._hit_min is not None:
            self._hit_min.fill_(1)

        def copy_in(c):
            g = self.inst[c]
            if id(g) in last_done:
                self.s_in.wait_event(last_done[id(g)])
            with torch.cuda.stream(self.s_in), torch.no_grad():
                for dst, src in zip(g.inputs, host_in):
                    dst.copy_(self._slice(src, c), non_blocking=True)
                ev_in[c].record(self.s_in)

        # chunk c + d is copied in right after chunk c has been replayed: an instance is shared by
        # chunks of one size that are at least d apart, so its previous replay is already recorded
        for c in range(min(d, n)):
            copy_in(c)
        scalars = [None] * len(host_out)
        for c in range(n):
            g = self.inst[c]
            comp.wait_event(ev_in[c])
            if id(g) in last_out:
                comp.wait_event(last_out[id(g)])
            g.graph.replay()
            for t in g._pending:                    # hit counters are rewritten by the next replay
                torch.minimum(self._hit_min, t.min(), out=self._hit_min)
            done = torch.cuda.Event()
            done.record(comp)
            last_done[id(g)] = done
            if c + d < n:
                copy_in(c + d)
            self.s_out.wait_event(done)
            with torch.cuda.stream(self.s_out):
                for k, (dst, src) in enumerate(zip(host_out, g.outputs)):
                    src = src.detach()
                    if dst.dim() == 0:
                        scalars[k] = src.clone() if scalars[k] is None else scalars[k] + src
                        if c == n - 1:
                            dst.copy_(scalars[k], non_blocking=True)
                    else:
                        self._slice(dst, c).copy_(src, non_blocking=True)
                out = torch.cuda.Event()
                out.record(self.s_out)
                last_out[id(g)] = out
        self.s_out.synchronize()
        comp.wait_stream(self.s_out)
        if self._hit_min is not None and int(self._hit_min.item()) == 0:
            from . import ops
            raise RuntimeError(ops._NO_HIT_MSG)
