"""Developer probe: per-image plane-gradient error of the CUDA path vs the eager-CUDA oracle at config-2 size,
and one image against the CPU oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch
import nfi_oracle as O
from nfi_b200 import decoder, render, synth

B, H, W, S, res = int(sys.argv[1]) if len(sys.argv) > 1 else 16, 128, 128, 64, 256
planes = synth.make_planes(B, res=res, seed=1); fd = synth.make_field(planes, seed=3)
pal = synth.make_palette(B, 10, seed=4); cam, focal = synth.make_cameras(B, seed=2, perspective=True)
g = torch.Generator().manual_seed(99)
u_s = torch.rand(B, H, W, S, generator=g); u_p = torch.rand(B * H * W, S, generator=g)
target = torch.rand(B, H, W, 3, generator=g) * 2 - 1
keys = ("W0", "b0", "W1", "b1", "beta", "alpha")

def run_oracle(dev, sl):
    lv = [t[sl].clone().to(dev).requires_grad_() for t in (planes, pal, cam, focal)]
    w = {k: fd[k].to(dev) for k in keys}
    fp = O.FieldParams(w["W0"], w["b0"], w["W1"], w["b1"], 1.4, True, 10, w["beta"], w["alpha"])
    n = lv[0].shape[0]
    r0 = sl.start * H * W
    rgb, depth, mask = O.render(lv[0].view(n, 3, 32, res, res), lv[1], fp, H, W, lv[2], lv[3], None, None, S, False,
                                u_s[sl].to(dev), u_p[r0:r0 + n * H * W].to(dev))
    loss = ((rgb - target[sl].to(dev)) ** 2).sum() / (H * W * 3) if os.environ.get("MSE") else (rgb - target[sl].to(dev)).abs().sum() / (H * W * 3)
    return rgb.detach().cpu(), torch.autograd.grad(loss, lv)[0].cpu()

def run_ours(sl):
    dev = "cuda"
    lv = [t[sl].clone().to(dev).requires_grad_() for t in (planes, pal, cam, focal)]
    fld = decoder.Field(*[fd[k].to(dev) for k in keys], 1.4, 10, True)
    n = lv[0].shape[0]
    r0 = sl.start * H * W
    rgb, depth, mask = render.render_planes(lv[0], lv[1], fld, H, W, lv[2], lv[3], None, None, S, randomize=True,
                                            u_strat=u_s[sl].to(dev), u_pdf=u_p[r0:r0 + n * H * W].to(dev))
    loss = ((rgb - target[sl].to(dev)) ** 2).sum() / (H * W * 3) if os.environ.get("MSE") else (rgb - target[sl].to(dev)).abs().sum() / (H * W * 3)
    return rgb.detach().cpu(), torch.autograd.grad(loss, lv)[0].cpu()

full = slice(0, B)
rgb_o, g_o = run_oracle("cuda", full)
rgb_m, g_m = run_ours(full)
for b in range(B):
    e = float((g_m[b] - g_o[b]).abs().max() / g_o[b].abs().max())
    er = float((rgb_m[b] - rgb_o[b]).abs().max())
    print("image %2d  d_planes rel err vs eager-cuda %.2e   rgb abs err %.2e   |g|max %.3e" % (b, e, er, float(g_o[b].abs().max())))
worst = max(range(B), key=lambda b: float((g_m[b] - g_o[b]).abs().max() / g_o[b].abs().max()))
sl = slice(worst, worst + 1)
_, g_cpu = run_oracle("cpu", sl)
_, g_one = run_ours(sl)
_, g_ocu = run_oracle("cuda", sl)
den = g_cpu[0].abs().max()
print("worst image %d alone:  ours vs cpu %.2e   eager-cuda(alone) vs cpu %.2e   eager-cuda(batch) vs cpu %.2e   ours(batch) vs cpu %.2e" % (
    worst, float((g_one[0] - g_cpu[0]).abs().max() / den), float((g_ocu[0] - g_cpu[0]).abs().max() / den),
    float((g_o[worst] - g_cpu[0]).abs().max() / den), float((g_m[worst] - g_cpu[0]).abs().max() / den)))
d = (g_m[worst] - g_o[worst]).abs()
idx = torch.nonzero(d == d.max())[0].tolist()
print("largest difference at (channel, y, x) =", idx, " ours", float(g_m[worst][tuple(idx)]), " eager-cuda", float(g_o[worst][tuple(idx)]),
      " cpu", float(g_cpu[0][tuple(idx)]))

def texel_report(name, a, b):
    dd = (a - b).abs().view(3, 32, res, res).amax(dim=1)              # [plane, y, x]
    bad = dd > 1e-3 * float(den)
    print("%s: texels off by > 1e-3 |g|max: %d (planes %s)" % (name, int(bad.sum()), [int(bad[p].sum()) for p in range(3)]))
    flat = dd.flatten().topk(6)
    for v, i in zip(flat.values.tolist(), flat.indices.tolist()):
        p, y, x = i // (res * res), (i // res) % res, i % res
        print("   plane %d (y %3d, x %3d)  diff %.2e" % (p, y, x, v / float(den)))

texel_report("ours(alone) vs cpu", g_one[0], g_cpu[0])
texel_report("eager-cuda(alone) vs cpu", g_ocu[0], g_cpu[0])
texel_report("ours(alone) vs eager-cuda(alone)", g_one[0], g_ocu[0])
