import os, sys, torch, numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import nfi_b200
from nfi_b200 import ops, decoder
from golden_io import load
g = load("decoder_sdf_attn")
cu = lambda a: torch.as_tensor(np.asarray(a)).cuda()
planes = cu(g["planes"]).requires_grad_(); pts = cu(g["pts"])
pcl = ops.planes_to_channels_last(planes)
cfg = ops.FieldConfig(10, False, 1.0)
wl = [cu(g[k]).clone().requires_grad_() for k in ("W0", "b0", "W1", "b1")]
y, _ = ops.field_on_points(pcl, None, *wl, None, None, pts / 1.4, cfg, mode=1)
gw = torch.autograd.grad((y * cu(g["gy"])).sum(), wl)
for a, k in zip(gw, ("dy_W0", "dy_b0", "dy_W1", "dy_b1")):
    ref = torch.as_tensor(np.asarray(g[k]))
    e = (a.cpu() - ref).abs()
    print(k, tuple(ref.shape), "max err %.3e  ref max %.3e" % (e.max(), ref.abs().max()))
    if e.dim() == 2:
        print("  err by row:", [round(float(x), 3) for x in e.max(dim=1).values[:64:4]])
        print("  err by col:", [round(float(x), 3) for x in e.max(dim=0).values[:32:2]])
print("pts shape", tuple(pts.shape))
