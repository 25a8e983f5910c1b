"""Small workload for compute-sanitizer (tools/sanitize.sh): every warp-specialised field-kernel
instantiation the bench and the tests use, at sizes a sanitizer run finishes in minutes."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from nfi_b200 import decoder, ops, render, synth

dev = "cuda"
B, H, W, S, res = 2, 16, 32, 64, 32
planes = synth.make_planes(B, res=res, seed=1).to(dev)
fd = synth.make_field(planes.cpu(), seed=3)
fld = decoder.Field(*[fd[k].to(dev) for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], 1.4, 10, True)
pal = synth.make_palette(B, 10, seed=4).to(dev)
cam, focal = synth.make_cameras(B, seed=2, perspective=True)
for no_cam in (False, True):                       # WANT_PTS = true / false instantiations of the backward
    for dt in (torch.float32, torch.bfloat16):     # fp32 / bf16 plane copies
        p = planes.clone().requires_grad_(); a = pal.clone().requires_grad_()
        c = cam.to(dev).requires_grad_(); f = focal.to(dev).requires_grad_()
        rgb, depth, mask = render.render_planes(p, a, fld, H, W, c, f, None, None, S, randomize=True,
                                                force_no_cam_grad=no_cam, plane_dtype=dt)
        (rgb.abs().mean() + mask.mean()).backward()
        torch.cuda.synchronize()
        print("render no_cam=%s %s ok" % (no_cam, dt), float(mask.mean()))
# explicit points, heads and raw-decoder modes (PTS = true instantiations)
pcl = ops.planes_to_channels_last(planes).detach().requires_grad_()
pts = (torch.rand(B, 300, 3, device=dev) * 2.8 - 1.4).requires_grad_()
for mode in (0, 1):
    y, _ = ops.field_on_points(pcl, pal if mode == 0 else None, *fld.tensors(), pts, fld.cfg, mode=mode)
    y.square().mean().backward()
    torch.cuda.synchronize()
    print("points mode", mode, "ok")
