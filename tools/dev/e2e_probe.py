import os, sys, time, torch
sys.path.insert(0, os.getcwd())
import bench
from nfi_b200 import _lib, decoder, render
from nfi_b200.host_pipeline import HostStepPipeline
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
B, H, W, S = 16, 128, 128, 64
planes_h, fd, pal_h, cam_h, focal_h, target_h = bench.make_inputs(B, H, W, S)
fld = decoder.Field(*[fd[k].to(dev) for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], bench.SCENE_RANGE, 10, True)
pin = lambda t: t.contiguous().pin_memory()
host_in = [pin(t) for t in (planes_h, pal_h, cam_h, focal_h, target_h)]
host_out = [pin(torch.empty(s)) for s in ((), (B, H, W, 3), (B, H, W), (B, H, W), tuple(planes_h.shape), tuple(pal_h.shape), (B, 4, 4), (B,))]
evs = []
def chunk_step(p, a, c, f, tg):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    p, a, c, f = [t.requires_grad_() for t in (p, a, c, f)]
    rgb, depth, mask = render.render_planes(p, a, fld, H, W, c, f, None, None, S, randomize=True, white_background=False, check_hits="defer")
    loss = (rgb - tg).abs().sum() * (1.0 / (H * W * 3))
    grads = torch.autograd.grad(loss, [p, a, c, f])
    e1.record(); evs.append((e0, e1))
    return (loss, rgb, depth, mask, *grads)
for n, outs in ((1, host_out), (4, host_out), (4, host_out[:4])):
    host_out_n = outs
    pipe = HostStepPipeline(chunk_step, B, n, dev)
    for _ in range(3): pipe.run(host_in, host_out_n)
    torch.cuda.synchronize(); evs.clear()
    t0 = time.perf_counter()
    for _ in range(5): pipe.run(host_in, host_out_n)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    comp = sum(a.elapsed_time(b) for a, b in evs) / 5
    print("chunks", n, "outs", len(host_out_n), "wall ms/step %.2f" % ((t1 - t0) * 1e3 / 5), "sum of chunk compute ms %.2f" % comp)
# raw copy bandwidth
x = host_in[0]; d = torch.empty_like(x, device=dev)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): d.copy_(x, non_blocking=True)
torch.cuda.synchronize(); t1 = time.perf_counter()
print("H2D GB/s %.1f" % (x.numel() * 4 * 5 / (t1 - t0) / 1e9))
y = host_out[4]
t0 = time.perf_counter()
for _ in range(5): y.copy_(d, non_blocking=True)
torch.cuda.synchronize(); t1 = time.perf_counter()
print("D2H GB/s %.1f" % (x.numel() * 4 * 5 / (t1 - t0) / 1e9))
