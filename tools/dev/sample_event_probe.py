"""Developer probe: which SAMPLE carries the isolated plane-gradient differences between the CPU oracle and the
eager-CUDA oracle at config-2 size (image `b` of the bench batch)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch
import nfi_oracle as O
from nfi_b200 import synth

b = int(sys.argv[1]) if len(sys.argv) > 1 else 15
B, H, W, S, res = 16, 128, 128, 64, 256
planes = synth.make_planes(B, res=res, seed=1); fd = synth.make_field(planes, seed=3)
pal = synth.make_palette(B, 10, seed=4); cam, focal = synth.make_cameras(B, seed=2, perspective=True)
g = torch.Generator().manual_seed(99)
u_s = torch.rand(B, H, W, S, generator=g); u_p = torch.rand(B * H * W, S, generator=g)
target = torch.rand(B, H, W, 3, generator=g) * 2 - 1
keys = ("W0", "b0", "W1", "b1", "beta", "alpha")
sl = slice(b, b + 1)
orig = O.field_eval

def run(dev):
    rec = []
    def patched(*a, **k):
        out = orig(*a, **k)
        out[0].retain_grad(); out[1].retain_grad()
        rec.append(out)
        return out
    O.field_eval = patched
    lv = [t[sl].clone().to(dev).requires_grad_() for t in (planes, pal, cam, focal)]
    w = {k: fd[k].to(dev) for k in keys}
    fp = O.FieldParams(w["W0"], w["b0"], w["W1"], w["b1"], 1.4, True, 10, w["beta"], w["alpha"])
    r0 = b * H * W
    rgb, depth, mask, aux = O.render(lv[0].view(1, 3, 32, res, res), lv[1], fp, H, W, lv[2], lv[3], None, None, S, False,
                                     u_s[sl].to(dev), u_p[r0:r0 + H * W].to(dev), return_aux=True)
    loss = ((rgb - target[sl].to(dev)) ** 2).sum() / (H * W * 3)
    loss.backward()
    O.field_eval = orig
    d = {"rgb": rgb.detach().cpu(), "gplanes": lv[0].grad.cpu()}
    for name, o in zip(("c", "f"), rec):
        d["sigma_" + name] = o[0].detach().cpu().view(H * W, S); d["gsigma_" + name] = o[0].grad.cpu().view(H * W, S)
        d["grgb_" + name] = o[1].grad.cpu().view(H * W, S, 3)
    d["z_c"] = aux["z_coarse"].detach().cpu().view(H * W, S); d["z_f"] = aux["z_fine"].detach().cpu().view(H * W, S)
    d["perm"] = aux["perm"].cpu().view(H * W, 2 * S)
    d["o"] = aux["origins"].detach().cpu().view(H * W, 3); d["d"] = aux["dirs"].detach().cpu().view(H * W, 3)
    return d

A, C = run("cpu"), run("cuda")
print("rgb max abs diff %.2e   perm differs on %d rays" % (float((A["rgb"] - C["rgb"]).abs().max()),
      int((A["perm"] != C["perm"]).any(dim=-1).sum())))
for k in ("gsigma_c", "gsigma_f", "grgb_c", "grgb_f"):
    dd = (A[k] - C[k]).abs()
    if dd.dim() == 3: dd = dd.amax(-1)
    v, i = dd.flatten().topk(3)
    print(k, "max |ref| %.3e" % float(A[k].abs().max()), [(int(j) // S, int(j) % S, "%.2e" % float(x)) for x, j in zip(v, i)])
dd = (A["gsigma_f"] - C["gsigma_f"]).abs() + (A["gsigma_c"] - C["gsigma_c"]).abs()
ray = int(dd.amax(-1).argmax())
print("ray", ray, "(h, w) =", divmod(ray, W))
for name in ("c", "f"):
    zs, i = (A["z_" + name][ray]).sort()
    print(name, "perm equal:", bool((A["perm"][ray] == C["perm"][ray]).all()))
torch.set_printoptions(precision=9, linewidth=200)
zall = torch.cat((A["z_c"][ray], A["z_f"][ray])); zallc = torch.cat((C["z_c"][ray], C["z_f"][ray]))
sg = torch.cat((A["sigma_c"][ray], A["sigma_f"][ray])); sgc = torch.cat((C["sigma_c"][ray], C["sigma_f"][ray]))
gs = torch.cat((A["gsigma_c"][ray], A["gsigma_f"][ray])); gsc = torch.cat((C["gsigma_c"][ray], C["gsigma_f"][ray]))
order = A["perm"][ray]
print("sorted pos | src | z cpu | z cuda | sigma cpu | sigma cuda | dL/dsigma cpu | dL/dsigma cuda")
for pos in range(2 * S):
    j = int(order[pos])
    if abs(float(gs[j] - gsc[j])) > 1e-3 * float(gs.abs().max()) or float(sg[j]) > 1.0:
        print("%3d %3d  %.7f %.7f  %.6e %.6e  %.6e %.6e%s" % (pos, j, float(zall[j]), float(zallc[j]), float(sg[j]), float(sgc[j]),
              float(gs[j]), float(gsc[j]), "   <-- order differs" if int(C["perm"][ray][pos]) != j else ""))

# where do the PLANE gradients differ, and which samples sit there?
gd = (A["gplanes"] - C["gplanes"]).abs().view(3, 32, res, res).amax(1)
den = float(A["gplanes"].abs().max())
tops = {}
for pl in range(3):
    v, i = gd[pl].flatten().max(0)
    tops[pl] = (int(i) // res, int(i) % res, float(v) / den)
print("largest plane-gradient differences (row, col, rel):", tops)
tx = torch.tensor([tops[0][1], tops[0][0], tops[1][0]], dtype=torch.float32)       # texel coords of the point (x0, x1, x2)
zz = torch.cat((A["z_c"], A["z_f"]), dim=1)                                            # [R, 128]
pts = A["o"][:, None, :] + A["d"][:, None, :] * zz[..., None]
tex = ((pts / 1.4 + 1) / 2 * (res - 1))
near = ((tex - tx).abs().amax(-1) < 1.5)
idx = torch.nonzero(near)
gs_a = torch.cat((A["gsigma_c"], A["gsigma_f"]), 1); gs_c = torch.cat((C["gsigma_c"], C["gsigma_f"]), 1)
gr_a = torch.cat((A["grgb_c"], A["grgb_f"]), 1); gr_c = torch.cat((C["grgb_c"], C["grgb_f"]), 1)
zc_all = torch.cat((C["z_c"], C["z_f"]), dim=1)
print("%d samples within 1.5 texels of (%s); those whose gradients differ:" % (len(idx), tx.tolist()))
for r, j in idx.tolist():
    da = abs(float(gs_a[r, j] - gs_c[r, j])); dr = float((gr_a[r, j] - gr_c[r, j]).abs().max())
    if da > 1e-9 or dr > 1e-8:
        print("  ray %5d sample %3d  z cpu %.7f cuda %.7f  tex %s  dsigma %.3e / %.3e  |drgb| %.3e / %.3e" % (
            r, j, float(zz[r, j]), float(zc_all[r, j]), [round(float(t), 3) for t in tex[r, j]], float(gs_a[r, j]), float(gs_c[r, j]),
            float(gr_a[r, j].abs().max()), float(gr_c[r, j].abs().max())))

# the same window from this repo's CUDA path
from nfi_b200 import decoder, render
dev = "cuda"
lv = [t[sl].clone().to(dev).requires_grad_() for t in (planes, pal, cam, focal)]
fld = decoder.Field(*[fd[k].to(dev) for k in keys], 1.4, 10, True)
r0 = b * H * W
rgb, depth, mask = render.render_planes(lv[0], lv[1], fld, H, W, lv[2], lv[3], None, None, S, randomize=True,
                                        u_strat=u_s[sl].to(dev), u_pdf=u_p[r0:r0 + H * W].to(dev))
loss = ((rgb - target[sl].to(dev)) ** 2).sum() / (H * W * 3)
g_ours = torch.autograd.grad(loss, lv)[0].cpu()
for pl in range(3):
    r, c, _ = tops[pl]
    ch = int((A["gplanes"] - C["gplanes"]).abs().view(3, 32, res, res)[pl, :, r, c].argmax())
    print("plane %d channel %d, rows %d..%d x cols %d..%d   (cpu | eager-cuda | ours), units of 1e-8" % (pl, ch, r - 1, r + 1, c - 1, c + 1))
    for rr in range(r - 1, r + 2):
        row = []
        for src in (A["gplanes"], C["gplanes"], g_ours):
            row.append(" ".join("%8.3f" % (float(x) * 1e8) for x in src.view(3, 32, res, res)[pl, ch, rr, c - 1:c + 2]))
        print("   " + "  |  ".join(row))
