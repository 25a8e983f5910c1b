"""Developer probe: the render step with the generator-training gradient set (decoder weights, beta,
alpha, planes, attention values) at config-2 size, next to the inversion gradient set of bench.py."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (sets up the nfi_b200 import path)
from nfi_b200 import _lib, decoder, render  # noqa: E402

dev = torch.device("cuda", 0)
B, H, W, S = int(os.environ.get("B", 16)), 128, 128, 64
planes_h, fd, pal_h, cam_h, focal_h, target_h = bench.make_inputs(B, H, W, S)
for train in (False, True):
    params = [fd[k].to(dev).requires_grad_(train) for k in ("W0", "b0", "W1", "b1", "beta", "alpha")]
    fld = decoder.Field(*params, bench.SCENE_RANGE, 10, True)
    planes, pal = planes_h.to(dev).requires_grad_(), pal_h.to(dev).requires_grad_()
    cam, focal, target = cam_h.to(dev), focal_h.to(dev), target_h.to(dev)
    wrt = [planes, pal] + (params if train else [])

    def step():
        rgb, depth, mask = render.render_planes(planes, pal, fld, H, W, cam, focal, None, None, S,
                                                randomize=True, white_background=False)
        loss = (rgb - target).abs().mean()
        return torch.autograd.grad(loss, wrt)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    _lib.PROFILE.start(timed=("nfi_field_fwd", "nfi_field_bwd"))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record()
    torch.cuda.synchronize()
    _lib.PROFILE.stop()
    durs = _lib.PROFILE.durations_ms()
    print("train grads" if train else "inversion grads (no pose)", "%.2f ms/step" % (e0.elapsed_time(e1) / 5),
          {k: round(sum(v) / len(v), 3) for k, v in durs.items() if "field" in k})
