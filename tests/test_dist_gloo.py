"""World-size-2 gloo tests (CPU) of the host-side sharding logic: shard bounds,
argument slicing as nn.DataParallel scatters them, the ragged report gather, max-over-ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nfi_b200 import dist as nd


def test_shard_range_partitions_the_batch():
    for batch in (0, 1, 3, 16, 32, 33):
        for world in (1, 2, 4, 8):
            spans = [nd.shard_range(batch, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        nd.shard_range(4, 2, 2)


def test_shard_batch_slices_per_image_tensors_only():
    args = dict(cam=torch.arange(32.).reshape(2, 4, 4).repeat(2, 1, 1), focal=torch.tensor(1.5),
                center=None, S=64, bbox=torch.zeros(4, 2, 2))
    out = nd.shard_batch(args, 4, rank=1, world=2)
    assert out["cam"].shape[0] == 2 and torch.equal(out["cam"], args["cam"][2:4])
    assert out["focal"].dim() == 0 and out["center"] is None and out["S"] == 64
    assert out["bbox"].shape[0] == 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, batch, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = nd.shard_range(batch)
        # each rank "renders" its own images: output row i of image b is b * 10 + i
        full = torch.arange(batch, dtype=torch.float32)[:, None] * 10 + torch.arange(3.)[None]
        local = full[lo:hi].clone()
        got = nd.gather_report(local, batch)
        ok = torch.equal(got, full)
        t = nd.max_over_ranks(1.0 + rank)
        lo2, hi2, out = nd.render_sharded(lambda scale, x: x * scale, batch, {"x": full}, shared=(2.0,))
        ok = ok and (lo2, hi2) == (lo, hi) and (out is None if hi == lo else torch.equal(out, full[lo:hi] * 2))
        # decoder-gradient bucket: every rank contributes (rank + 1) * pattern; None entries are skipped
        pats = [torch.arange(6.).view(2, 3), None, torch.ones(4)]
        grads = [None if p is None else p * (rank + 1) for p in pats]
        nd.allreduce_grads(grads)
        tot = sum(r + 1 for r in range(world))
        ok = ok and grads[1] is None and torch.equal(grads[0], pats[0] * tot) and torch.equal(grads[2], pats[2] * tot)
        nd.allreduce_grads(grads, batch_weight=1.0 / world)
        ok = ok and torch.allclose(grads[0], pats[0] * tot)      # mean of equal contributions
        # a rank that owns no image (batch < world) passes None for everything and must still enter the same
        # collectives: rank 1 contributes nothing, both ranks end with rank 0's gradients
        mine = [p.clone() if (p is not None and rank == 0) else None for p in pats]
        out = nd.allreduce_grads(mine)
        ok = ok and out[1] is None and torch.equal(out[0].reshape(2, 3), pats[0]) and torch.equal(out[2], pats[2])
        shapes = [torch.empty(2, 3), torch.empty(5), torch.empty(4)]
        mine = [pats[0].clone() if rank == 0 else None, None, pats[2].clone() if rank == 1 else None]
        out = nd.allreduce_grads(mine, params=shapes)
        ok = ok and out[0].shape == (2, 3) and torch.equal(out[0], pats[0]) and torch.equal(out[2], pats[2])
        ok = ok and out[1].shape == (5,) and float(out[1].abs().sum()) == 0.0
        q.put((rank, bool(ok), t))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("batch", [4, 5, 1])
def test_gloo_world2_gather_and_timing(batch):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[0] for r in res] == [0, 1]
    assert all(r[1] for r in res)
    assert all(r[2] == 2.0 for r in res)      # max over ranks of (1 + rank)
