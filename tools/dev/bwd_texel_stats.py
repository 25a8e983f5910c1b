"""Developer probe (CPU): how many DISTINCT plane-gradient texel lines does a 128-sample tile of the
config-2 workload touch, for (a) the current pixel-block x one-depth tiling, (b) ray-major tiles
(2 rays x 64 samples), (c) 8x16 pixel block x 64 depths (window).  Guides the in-SM aggregation design."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "oracle"))
import nfi_oracle as O
from nfi_b200 import synth

B, H, W, S, res, sr = 1, 128, 128, 64, 256, 1.4
cam, focal = synth.make_cameras(2, seed=2, perspective=True)
cam, focal = cam[:1], focal[:1]
o, d = O.pixel_rays(H, W, cam, focal); d = O.unit_directions(d)
o = o.expand_as(d)
near, far, hit = O.box_near_far(o.reshape(-1,3), d.reshape(-1,3), sr, return_hit=True)
near = near.reshape(H, W); far = far.reshape(H, W)
g = torch.Generator().manual_seed(0)
u = torch.rand(H, W, S, generator=g)
t = (torch.arange(S) + u) / S
z = near[..., None] + (far - near)[..., None] * t          # coarse
# fine: concentrate around a surface at 40-60% depth with small spread
zc = near + (far - near) * (0.45 + 0.1 * torch.rand(H, W, generator=g))
zf = (zc[..., None] + 0.03 * torch.randn(H, W, S, generator=g)).clamp(min=near[..., None], max=far[..., None])
def texels(z):
    pts = o[0][..., None, :] + d[0][..., None, :] * z[..., None]       # H,W,S,3
    x = pts / sr
    inb = (x.abs() <= 1).all(-1)
    out = []
    for (a, b_) in ((0, 1), (0, 2), (1, 2)):
        ix = ((x[..., a] + 1) * res - 1) / 2; iy = ((x[..., b_] + 1) * res - 1) / 2
        x0 = ix.floor().clamp(0, res - 1).long(); y0 = iy.floor().clamp(0, res - 1).long()
        x1 = (x0 + 1).clamp(max=res - 1); y1 = (y0 + 1).clamp(max=res - 1)
        out.append(torch.stack((y0 * res + x0, y0 * res + x1, y1 * res + x0, y1 * res + x1), -1))
    return torch.stack(out, -2), inb        # H,W,S,3,4
for name, zz in (("coarse", z), ("fine", zf)):
    tx, inb = texels(zz)
    tx = tx + torch.arange(3).view(1, 1, 1, 3, 1) * res * res
    tx = torch.where(inb[..., None, None], tx, torch.full_like(tx, -1))
    def distinct(groups):
        tot = 0; n = 0; act = 0
        for grp in groups:
            v = grp.reshape(-1); v = v[v >= 0]
            if v.numel() == 0: continue
            tot += torch.unique(v).numel(); n += 1; act += v.numel()
        return tot, n, act
    # (a) 8x16 block at one depth
    ga = [tx[y:y+8, x:x+16, s] for y in range(0, H, 32) for x in range(0, W, 32) for s in range(0, S, 4)]
    gb = [tx[y, x:x+2, :] for y in range(0, H, 16) for x in range(0, W, 16)]
    gc = [tx[y:y+8, x:x+16, :] for y in range(0, H, 32) for x in range(0, W, 32)]
    gd = [tx[y:y+4, x:x+4, s:s+8] for y in range(0, H, 32) for x in range(0, W, 32) for s in range(0, S, 16)]
    ge = [tx[y:y+8, x:x+16, s:s+8] for y in range(0, H, 32) for x in range(0, W, 32) for s in range(0, S, 16)]
    for nm, gg in (("8x16 x1 depth (128)", ga), ("2 rays x 64 (128)", gb), ("8x16 x 64 depths (8192)", gc), ("4x4 x 8 depths (128)", gd), ("8x16 x 8 depths (1024)", ge)):
        tot, n, act = distinct(gg)
        print("%-7s %-26s groups %4d  red lines now %8d  distinct %8d  ratio %.2f" % (name, nm, n, act, tot, act / max(tot, 1)))
