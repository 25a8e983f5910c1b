"""Does a D2H / H2D copy on another stream slow a compute-stream kernel sequence?"""
import os, sys, time, torch
sys.path.insert(0, os.getcwd())
import bench
from nfi_b200 import decoder, render
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
B, H, W, S = 4, 128, 128, 64
planes_h, fd, pal_h, cam_h, focal_h, target_h = bench.make_inputs(B, H, W, S)
fld = decoder.Field(*[fd[k].to(dev) for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], bench.SCENE_RANGE, 10, True)
leaves = [t.to(dev).requires_grad_() for t in (planes_h, pal_h, cam_h, focal_h)]
tg = target_h.to(dev)
def step():
    rgb, depth, mask = render.render_planes(*leaves[:2], fld, H, W, leaves[2], leaves[3], None, None, S, randomize=True, white_background=False, check_hits="defer")
    loss = (rgb - tg).abs().sum()
    return torch.autograd.grad(loss, leaves)
big_d = torch.empty(100 * 1024 * 1024 // 4, device=dev)
big_h = torch.empty(100 * 1024 * 1024 // 4).pin_memory()
side = torch.cuda.Stream()
def timed(copy):
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(8):
        if copy == "d2h":
            with torch.cuda.stream(side): big_h.copy_(big_d, non_blocking=True)
        elif copy == "h2d":
            with torch.cuda.stream(side): big_d.copy_(big_h, non_blocking=True)
        step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 8
for c in ("none", "h2d", "d2h", "none"):
    print(c, "ms/step %.2f" % timed(c))
