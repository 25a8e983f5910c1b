"""Developer stress test of the host pipelines: many runs of bench.py's end-to-end chunk pipeline,
checking every run's outputs (finite, loss in range).  MODE=graph|eager, RUNS=n, CHUNKS=1,3,4,4,3,1"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from nfi_b200 import decoder, render  # noqa: E402
from nfi_b200.host_pipeline import GraphedHostStepPipeline, HostStepPipeline  # noqa: E402

dev = torch.device("cuda", 0)
B, H, W, S = 16, 128, 128, 64
mode = os.environ.get("MODE", "graph")
runs = int(os.environ.get("RUNS", 200))
chunks = [int(v) for v in os.environ.get("CHUNKS", "1,3,4,4,3,1").split(",")]
planes_h, fd, pal_h, cam_h, focal_h, target_h = bench.make_inputs(B, H, W, S)
fld = decoder.Field(*[fd[k].to(dev) for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], bench.SCENE_RANGE, 10, True)
pin = lambda t: t.contiguous().pin_memory()
host_in = [pin(t) for t in (planes_h, pal_h, cam_h, focal_h, target_h)]
host_out = [pin(torch.empty(s)) for s in ((), (B, H, W, 3), (B, H, W), (B, H, W), tuple(planes_h.shape),
                                          tuple(pal_h.shape), (B, 4, 4), (B,))]


def chunk_step(p, a, c, f, tg):
    p, a, c, f = [t.requires_grad_() for t in (p, a, c, f)]
    rgb, depth, mask = render.render_planes(p, a, fld, H, W, c, f, None, None, S, randomize=True,
                                            white_background=False, check_hits="defer")
    loss = (rgb - tg).abs().sum() * (1.0 / (H * W * 3))
    return (loss, rgb, depth, mask, *torch.autograd.grad(loss, [p, a, c, f]))


pipe = GraphedHostStepPipeline(chunk_step, B, chunks, dev, host_in) if mode == "graph" \
    else HostStepPipeline(chunk_step, B, chunks, dev)
losses = []
t0 = time.time()
for i in range(runs):
    pipe.run(host_in, host_out)
    assert all(bool(torch.isfinite(t).all()) for t in host_out), "non-finite output in run %d" % i
    losses.append(float(host_out[0]))
    if i and abs(losses[-1] - losses[0]) > 0.05 * abs(losses[0]):
        raise SystemExit("run %d: loss %.6f far from first %.6f" % (i, losses[-1], losses[0]))
print("%s: %d runs ok in %.1f s, loss %.5f .. %.5f" % (mode, runs, time.time() - t0, min(losses), max(losses)))
