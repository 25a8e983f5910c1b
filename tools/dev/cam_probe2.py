import os, sys, torch
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "oracle"))
import nfi_oracle as O
from nfi_b200 import ops, decoder, render, synth
B, H, W, S, res, seed = 1, 32, 32, 64, 64, 0
planes = synth.make_planes(B, res=res, seed=1 + seed); fd = synth.make_field(planes, seed=3 + seed)
pal = synth.make_palette(B, 10, seed=4 + seed); cam, focal = synth.make_cameras(B, seed=2 + seed, perspective=True)
g = torch.Generator().manual_seed(99 + seed)
u_s = torch.rand(B, H, W, S, generator=g); u_p = torch.rand(B * H * W, S, generator=g); target = torch.rand(B, H, W, 3, generator=g) * 2 - 1
wk = ("W0", "b0", "W1", "b1", "beta", "alpha")
def run(dev):
    p = planes.clone().to(dev).requires_grad_(); a = pal.clone().to(dev); c = cam.clone().to(dev).requires_grad_(); f = focal.clone().to(dev).requires_grad_()
    if dev == "cpu":
        fp = O.FieldParams(*[fd[k] for k in wk[:4]], 1.4, True, 10, fd["beta"], fd["alpha"])
        rgb, depth, mask, aux = O.render(p.view(B, 3, 32, res, res), a, fp, H, W, c, f, None, None, S, False, u_s, u_p, True, True, False, return_aux=True)
    else:
        fld = decoder.Field(*[fd[k].to(dev) for k in wk], 1.4, 10, True)
        rgb, depth, mask, aux = render.render_planes(p, a, fld, H, W, c, f, None, None, S, randomize=True, white_background=False, u_strat=u_s.to(dev), u_pdf=u_p.to(dev), return_aux=True)
    loss = (rgb - target.to(dev)).abs().mean() * B + 0.1 * mask.mean()
    go, gd, gc, gf = torch.autograd.grad(loss, [aux["origins"], aux["dirs"], c, f])
    return go.detach().cpu().reshape(-1, 3), gd.detach().cpu().reshape(-1, 3), gc.detach().cpu(), gf.detach().cpu()
r = run("cpu"); q = run("cuda")
for name, a, b in (("d_origins", q[0], r[0]), ("d_dirs", q[1], r[1])):
    e = (a - b).abs().max(dim=1).values
    print(name, "max ref %.3e" % b.abs().max(), "max err %.3e" % e.max(), "worst rays", e.topk(5).indices.tolist(), ["%.2e" % x for x in e.topk(5).values.tolist()])
    print("   sum over rays: cuda", a.sum(0).tolist(), "ref", b.sum(0).tolist())
print("d_cam", q[2].flatten().tolist(), "\n ref ", r[2].flatten().tolist())
print("d_focal", q[3].tolist(), r[3].tolist())
# which samples of the worst rays sit on a discontinuity?
p = planes.clone(); c = cam.clone(); f = focal.clone()
fp = O.FieldParams(*[fd[k] for k in wk[:4]], 1.4, True, 10, fd["beta"], fd["alpha"])
rgb, depth, mask, aux = O.render(p.view(B, 3, 32, res, res), pal, fp, H, W, c, f, None, None, S, False, u_s, u_p, True, True, False, return_aux=True)
o = aux["origins"].reshape(-1, 3); d = aux["dirs"].reshape(-1, 3)
for ray in (144, 643, 695, 100):
    for nm, z in (("coarse", aux["z_coarse"].reshape(-1, S)[ray]), ("fine", aux["z_fine"].reshape(-1, S)[ray])):
        pts = o[ray][None] + d[ray][None] * z[:, None]
        x = pts / 1.4
        ix = ((x + 1) / 2) * (res - 1)
        frac = (ix - ix.round()).abs()
        edge = (x.abs() - 1).abs()
        print(ray, nm, "min |ix - int| %.2e" % frac.min(), "min ||x|-1| %.2e" % edge.min(), "n in box", int((x.abs() <= 1).all(dim=1).sum()))
