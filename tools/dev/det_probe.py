"""Developer probe (GPU): deterministic-mode render, ours vs the lifted reference render on the same device."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import ref_loader, ref_generator as rg
from nfi_b200 import render as R, synth
torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False
G = rg.build_generator("cuda")
B, HW = 2, 32
ws = rg.make_ws(B, "cuda")
cam, focal = synth.make_cameras(B, seed=2, perspective=True); cam = cam.cuda(); focal = focal.cuda()
ref_render, ns = ref_loader.lift_render("run.py", scene_range=1.4)
nu = ns["nerf_utils"]
rec = {}
orig_pdf = nu.sample_pdf
def spy_pdf(bins, w, n, deterministic=False):
    out = orig_pdf(bins, w, n, deterministic)
    rec["bins"], rec["w"], rec["zf"] = bins.clone(), w.clone(), out.clone()
    return out
class NS: pass
fake = NS(); fake.__dict__.update({k: getattr(nu, k) for k in dir(nu) if not k.startswith("__")}); fake.sample_pdf = spy_pdf
ns["nerf_utils"] = fake
with torch.no_grad():
    ref = ref_render(G, HW, HW, cam, focal, None, None, ws, 64, randomize=False, extra_model_inputs={"freeze_noise": True})
    mo = G(None, ws, ["sampler"], {"freeze_noise": True}); cv = R._closure_vars(mo["sampler"])
    from nfi_b200.decoder import Field, _as_planes_tensor
    planes = _as_planes_tensor(cv["xy"], cv["xz"], cv["yz"])
    out = R.render_planes(planes, cv["attention_values"], Field.from_generator(G), HW, HW, cam, focal, None, None, 64,
                          randomize=False, return_aux=True)
rgb, depth, mask, aux = out
d = (rgb - ref[0]).abs().amax(-1)
i = int(d.argmax()); print("max rgb diff", float(d.max()), "at ray", i, "rgb scale", float(ref[0].abs().max()))
zf_ref = rec["zf"].reshape(-1, 64); zf = aux["z_fine"].reshape(-1, 64)
dz = (zf - zf_ref).abs()
print("z_fine max diff", float(dz.max()), "rays with diff>1e-4:", int((dz.amax(-1) > 1e-4).sum()), "of", dz.shape[0])
print("z_coarse diff", float((aux["z_coarse"].reshape(-1, 64) - 0).abs().max()))
r = int(dz.amax(-1).argmax())
print("ray", r, "ours", zf[r][:8].tolist(), zf[r][-4:].tolist()); print("ref ", zf_ref[r][:8].tolist(), zf_ref[r][-4:].tolist())
k = int(dz[r].argmax()); print("sample", k, float(zf[r, k]), float(zf_ref[r, k]))
w = rec["w"][r] + 1e-5; pdf = w / w.sum(); cdf = torch.cumsum(pdf, -1)
print("cdf tail", cdf[-3:].tolist(), "u_k", k / 63.0)
ws_ = torch.sort(dz.flatten(), descending=True)[0][:10]; print("top dz", ws_.tolist())
print("ray of max rgb diff: dz", float(dz[i].max()))
# --- are the knife-edge rays explained by the last bit of the normalised directions?
import torch.nn.functional as F
o_r, d_r = nu.get_ray_bundle(HW, HW, focal, cam, None, None)
d_r = F.normalize(d_r, dim=-1)
d_cpu = F.normalize(nu.get_ray_bundle(HW, HW, focal.cpu(), cam.cpu(), None, None)[1], dim=-1)
print("dirs: CUDA ATen vs ours bit-equal:", bool((d_r == aux["dirs"]).all()), " CPU ATen vs ours:", bool((d_cpu == aux["dirs"].cpu()).all()),
      " CUDA vs CPU ATen differing elements:", int((d_r.cpu() != d_cpu).sum()), "of", d_cpu.numel())
bad = (dz.amax(-1) > 1e-4)
n_r, f_r = nu.compute_near_far_planes(o_r.expand_as(d_r).contiguous(), d_r, 1.4)
print("near bit-equal:", bool((n_r == aux["near"]).all()), "max near diff", float((n_r - aux["near"]).abs().max()))
p0_ref = (o_r.expand_as(d_r) + d_r * n_r[..., None]).reshape(-1, 3) / 1.4
p0_our = (aux["origins"] + aux["dirs"] * aux["near"][..., None]).reshape(-1, 3) / 1.4
in_ref = ~(p0_ref.abs() > 1).any(-1); in_our = ~(p0_our.abs() > 1).any(-1)
print("first sample inside box: ref", int(in_ref.sum()), "ours", int(in_our.sum()), " disagree on", int((in_ref != in_our).sum()),
      "rays; of the", int(bad.sum()), "rays with different z_fine,", int((bad & (in_ref != in_our)).sum()), "are among them")
