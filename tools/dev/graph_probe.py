"""Developer probe: eager vs CUDA-graph replay of the config-2 step (time per step, same results)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
from nfi_b200 import decoder, render, graph

dev = torch.device("cuda:0")
B, H, W, S = int(sys.argv[1]) if len(sys.argv) > 1 else 16, 128, 128, 64
planes_h, fd, pal_h, cam_h, focal_h, target_h = bench.make_inputs(B, H, W, S)
fld = decoder.Field(*[fd[k].to(dev) for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], 1.4, 10, True)
u_s = torch.rand(B, H, W, S, device=dev); u_p = torch.rand(B * H * W, S, device=dev)

def make(fixed_u):
    def step(planes, pal, cam, focal, target):
        rgb, depth, mask = render.render_planes(planes, pal, fld, H, W, cam, focal, None, None, S, randomize=True,
                                                white_background=False, check_hits="defer",
                                                u_strat=u_s if fixed_u else None, u_pdf=u_p if fixed_u else None)
        loss = (rgb - target).abs().mean() * B
        return (loss, rgb, depth, mask, *torch.autograd.grad(loss, [planes, pal, cam, focal]))
    return step

leaves = [t.to(dev).requires_grad_() for t in (planes_h, pal_h, cam_h, focal_h)] + [target_h.to(dev)]
def timeit(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

eager = make(False)
print("eager   ms/step", timeit(lambda: eager(*leaves)))
g = graph.GraphedStep(make(False), leaves)
print("graphed ms/step", timeit(lambda: g()), "launches/replay", g.launches_per_replay)
g.check_hits()
# same results with fixed uniforms
ref = make(True)(*leaves)
g2 = graph.GraphedStep(make(True), leaves)
out = g2()
torch.cuda.synchronize()
for name, a, b in zip(("loss", "rgb", "depth", "mask", "d_planes", "d_pal", "d_cam", "d_focal"), out, ref):
    print(name, float((a - b).abs().max()), float(b.abs().max()))
# fresh randomness per replay
l1 = float(g()[0]); l2 = float(g()[0]); print("losses of two replays", l1, l2)
