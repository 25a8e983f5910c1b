"""Per-image sharding of the render path across ranks.

The path has no exchange step: every image -- planes, palette, pose, rays, samples, outputs and
gradients -- is independent (run.py scatters along dim 0 through nn.DataParallel, run.py:636-644).
Rank r of G owns images [r*B/G, (r+1)*B/G); there is NO collective on the data path.  The helpers
here are the host-side plumbing around that: shard bounds, slicing a batch of render arguments, and
the one collective the inversion loop has -- gathering per-image report tensors at the end
(run.py:2020-2027) -- plus the gradient all-reduce of generator training for the parameters that live
on this path (decoder weights, beta, alpha; the reference gets it implicitly from nn.DataParallel's
gradient reduction, run.py:636-644).  Works on any torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests).
"""
from typing import Any, Dict, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def rank_world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(batch: int, rank: Optional[int] = None, world: Optional[int] = None) -> Tuple[int, int]:
    """Images [lo, hi) owned by `rank`.  Ragged batches are allowed: the first batch % world ranks
    get one extra image; a rank may own nothing (lo == hi) when batch < world."""
    if rank is None or world is None:
        rank, world = rank_world()
    if batch < 0 or world <= 0 or not 0 <= rank < world:
        raise ValueError("shard_range: bad batch/rank/world = %r/%r/%r" % (batch, rank, world))
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(args: Dict[str, Any], batch: int, rank: Optional[int] = None,
                world: Optional[int] = None) -> Dict[str, Any]:
    """Slice every per-image tensor of a dict of render arguments (leading dim == batch) to this
    rank's images; 0-dim tensors (the perspective-eval focal, eval_kitti_persp.py:948), None and
    non-tensors are shared by all ranks, as nn.DataParallel's scatter does."""
    lo, hi = shard_range(batch, rank, world)
    out = {}
    for k, v in args.items():
        if torch.is_tensor(v) and v.dim() > 0 and v.shape[0] == batch:
            out[k] = v[lo:hi]
        else:
            out[k] = v
    return out


def gather_report(local: torch.Tensor, batch: int) -> torch.Tensor:
    """all_gather of per-image report tensors (ws, pose parameters, PSNR ...; run.py:2020-2027) back
    into batch order.  `local` holds this rank's images [hi-lo, ...]; shards may be ragged."""
    rank, world = rank_world()
    if world == 1:
        return local
    lo, hi = shard_range(batch, rank, world)
    if local.shape[0] != hi - lo:
        raise ValueError("gather_report: rank %d owns %d images, got %d" % (rank, hi - lo, local.shape[0]))
    width = max(shard_range(batch, r, world)[1] - shard_range(batch, r, world)[0] for r in range(world))
    pad = local.new_zeros((width,) + tuple(local.shape[1:]))
    pad[: hi - lo] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad.contiguous())
    sizes = [shard_range(batch, r, world) for r in range(world)]
    return torch.cat([p[: b - a] for p, (a, b) in zip(parts, sizes)], dim=0)


def _collective_device(tensors) -> torch.device:
    for t in tensors:
        if t is not None:
            return t.device
    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def allreduce_grads(grads: Sequence[Optional[torch.Tensor]], batch_weight: Optional[float] = None,
                    params: Optional[Sequence[torch.Tensor]] = None):
    """Sum the gradients of parameters shared by all images (decoder W0/b0/W1/b1, beta, alpha) over the
    ranks as ONE flat bucket (they are ~2.7 k floats in total: one collective of launch-latency size, not
    one per tensor).  nn.DataParallel sums replica gradients (run.py:636-644), so the default is a plain
    sum; `batch_weight` scales the result (e.g. 1 / world for a mean).

    Every rank always enters the same two collectives with buffers of identical length, whatever it holds:
    a rank that owns no image (shard_range allows batch < world) passes None for everything and contributes
    zeros.  The bucket layout comes from `params` (same list on every rank) when given, else from a MAX
    all-reduce of the per-entry element counts.  Returns a list: live entries are reduced in place, entries
    that were None here but live on another rank come back as new tensors (shaped like `params[i]`, or flat),
    entries that are None on every rank stay None."""
    rank, world = rank_world()
    grads = list(grads)
    if params is not None and len(params) != len(grads):
        raise ValueError("allreduce_grads: %d gradients for %d parameters" % (len(grads), len(params)))
    if world == 1:
        if batch_weight is not None:
            for g in grads:
                if g is not None:
                    g.mul_(batch_weight)
        return grads
    dev = _collective_device(list(grads) + list(params or []))
    if params is not None:
        sizes = [int(p.numel()) for p in params]
    else:
        t = torch.tensor([0 if g is None else g.numel() for g in grads], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sizes = [int(v) for v in t.tolist()]
    for g, n in zip(grads, sizes):
        if g is not None and g.numel() != n:
            raise ValueError("allreduce_grads: a gradient has %d elements here, %d elsewhere" % (g.numel(), n))
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
    off = 0
    for g, n in zip(grads, sizes):
        if g is not None:
            flat[off:off + n] = g.reshape(-1).to(torch.float32)
        off += n
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    if batch_weight is not None:
        flat.mul_(batch_weight)
    off = 0
    for i, (g, n) in enumerate(zip(grads, sizes)):
        if n == 0:
            continue
        if g is not None:
            g.copy_(flat[off:off + n].view_as(g))
        else:
            piece = flat[off:off + n].clone()
            grads[i] = piece.view_as(params[i]) if params is not None else piece
        off += n
    return grads


def max_over_ranks(value: float, device: Optional[torch.device] = None) -> float:
    """Timing rule of bench.py: a multi-GPU number is the max over ranks."""
    rank, world = rank_world()
    if world == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def render_sharded(render_fn, batch: int, per_image: Dict[str, Any], shared: Sequence[Any] = ()):
    """Run `render_fn(*shared, **shard)` on this rank's images only and return (lo, hi, outputs).
    No collective: the caller keeps its slice (inversion) or reduces its own loss."""
    lo, hi = shard_range(batch)
    if hi == lo:
        return lo, hi, None
    return lo, hi, render_fn(*shared, **shard_batch(per_image, batch))
