"""Generate tests/golden/*.npz by running the UNMODIFIED reference -- TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference):

    PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden.py

Every array stored as an *output* below was produced by the reference's own code
(`lib/nerf_utils.py`, `models/generator.py`, `run.py::render` / `eval_kitti_persp.py::render`
lifted with `ast`); inputs come from the seeded generators in nfi_b200.synth.  The fixtures
travel to the GPU box, where /root/reference does not exist.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path[:0] = [ROOT, HERE]

import ref_loader  # noqa: E402
from nfi_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def npy(d):
    return {k: (v.detach().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()
            if v is not None}


class PlanesStub(torch.nn.Module):
    def __init__(self, planes):
        super().__init__()
        self.planes = planes

    def forward(self, ws, **kw):
        return self.planes


def reference_generator(planes, field, scene_range, A=10, use_sdf=True):
    gen_mod = ref_loader.modules().generator
    torch.manual_seed(0)
    G = gen_mod.Generator(512, scene_range, attention_values=A, use_viewdir=False, use_encoder=False,
                          disable_stylegan_noise=True, use_sdf=use_sdf).eval()
    G.synthesis_network = PlanesStub(planes)
    with torch.no_grad():
        G.decoder.net[0].weight.copy_(field["W0"]); G.decoder.net[0].bias.copy_(field["b0"])
        G.decoder.net[2].weight.copy_(field["W1"]); G.decoder.net[2].bias.copy_(field["b1"])
        if use_sdf:
            G.beta.copy_(field["beta"]); G.alpha.copy_(field["alpha"])
    return G.requires_grad_(False)


def golden_functions():
    """Per-function vectors for lib/nerf_utils.py (A1-A3, A7, A8, A10)."""
    nu = ref_loader.modules().nerf_utils
    g = torch.Generator().manual_seed(11)
    out = {}
    H, W, B = 10, 14, 3
    cam_p, focal = synth.make_cameras(B, seed=2)
    cam_o, _ = synth.make_cameras(B, seed=5, perspective=False)
    cam_c = synth.make_ortho_cub_cameras(B, seed=7)
    bbox = torch.stack((torch.rand(B, 2, generator=g) - 0.5, 0.5 + torch.rand(B, 2, generator=g)), 1)
    center = 0.5 + (torch.rand(B, 2, generator=g) - 0.5) * 0.6
    out.update(HW=np.array([H, W]), cam_persp=cam_p, focal=focal, cam_ortho=cam_o, cam_cub=cam_c,
               bbox=bbox, center=center)
    variants = {"persp": (focal, cam_p, None, None), "persp_c": (focal, cam_p, None, center),
                "persp_b": (focal, cam_p, bbox, None), "persp_bc": (focal, cam_p, bbox, center),
                "ortho": (None, cam_o, None, None), "ortho_b": (None, cam_o, bbox, None),
                "cub": (None, cam_c, None, None)}
    for name, (f, cam, bb, c) in variants.items():
        o, d = nu.get_ray_bundle(H, W, f, cam, bb, c)
        out["rays_o_" + name], out["rays_d_" + name] = o.expand_as(d).contiguous(), d
    o, d = nu.get_ray_bundle(H, W, focal[0].clone(), cam_p[:1], None, center[:1])   # 0-dim focal
    out["rays_d_persp_scalar_focal"] = d
    # near/far: a small box so that some rays miss
    for name, r in (("persp", 1.0), ("ortho", 0.6), ("cub", 2.0)):
        f, cam, _, _ = variants[name]
        o, d = nu.get_ray_bundle(H, W, f, cam, None, None)
        d = F.normalize(d, dim=-1)
        n, fa = nu.compute_near_far_planes(o, d, r)
        out["near_" + name], out["far_" + name] = n, fa
        out["unit_d_" + name] = d
        if name == "persp":
            torch.manual_seed(5)
            q, z = nu.compute_query_points_from_rays(o, d, n, fa, 64, randomize=True)
            torch.manual_seed(5)
            out["strat_u"] = torch.rand(B, H, W, 64)
            out["strat_z_rand"], out["strat_pts_rand"] = z, q
            q, z = nu.compute_query_points_from_rays(o, d, n, fa, 64, randomize=False)
            out["strat_z_det"] = z
    # weights / smoothing / sample_pdf / composite on random rays
    R, S = 256, 64
    dd = F.normalize(torch.randn(R, 3, generator=g), dim=-1) * (0.5 + torch.rand(R, 1, generator=g))
    near = 0.5 + torch.rand(R, generator=g)
    far = near + 0.5 + 2 * torch.rand(R, generator=g)
    z = torch.lerp(near[:, None], far[:, None], torch.arange(S) / S)
    z = z + torch.rand(R, S, generator=g) * ((far - near) / S).unsqueeze(-1)
    sigma = torch.rand(R, S, generator=g) * (torch.rand(R, S, generator=g) > 0.6) * 30
    w = nu.render_volume_density_weights_only(sigma, torch.zeros(R, 3), dd, z)
    ws = F.avg_pool1d(F.max_pool1d(w.unsqueeze(1), 2, 1, padding=1), 2, 1).squeeze() + 0.01
    mid = 0.5 * (z[..., 1:] + z[..., :-1])
    torch.manual_seed(9)
    zs_r = nu.sample_pdf(mid, ws[..., 1:-1], S, deterministic=False)
    torch.manual_seed(9)
    u = torch.rand(R, S)
    zs_d = nu.sample_pdf(mid, ws[..., 1:-1], S, deterministic=True)
    out.update(w_dirs=dd, w_z=z, w_sigma=sigma, w_weights=w, w_smooth=ws, pdf_u=u,
               pdf_z_rand=zs_r, pdf_z_det=zs_d)
    S2, R = 128, 96
    dd = dd[:R]
    z2 = torch.sort(torch.rand(R, S2, generator=g) * 3 + 0.2, dim=-1).values
    sigma2 = torch.rand(R, S2, generator=g) * 20 * (torch.rand(R, S2, generator=g) > 0.3)
    rgb2 = torch.rand(R, S2, 3, generator=g) * 2 - 1
    sem2 = torch.rand(R, S2, 10, generator=g)
    nrm2 = F.normalize(torch.randn(R, S2, 3, generator=g), dim=-1)
    out.update(c_dirs=dd, c_z=z2, c_sigma=sigma2, c_rgb=rgb2, c_sem=sem2, c_nrm=nrm2)
    for wb in (0, 1):
        sig = sigma2.clone().requires_grad_(); col = rgb2.clone().requires_grad_()
        ddg = dd.clone().requires_grad_()
        res = nu.render_volume_density(sig, col, torch.zeros(R, 3), ddg, z2, nrm2, sem2,
                                       white_background=bool(wb))
        for k, v in zip(("rgb", "depth", "mask", "nrm", "sem"), res):
            out["c%d_%s" % (wb, k)] = v
        gr = torch.rand(R, 3, generator=torch.Generator().manual_seed(21)) - 0.5
        gm = torch.rand(R, generator=torch.Generator().manual_seed(22)) - 0.5
        gs, gc, gd = torch.autograd.grad((res[0] * gr).sum() + (res[2] * gm).sum(), (sig, col, ddg))
        out.update({"c_grad_rgb": gr, "c_grad_mask": gm, "c%d_dsigma" % wb: gs, "c%d_drgb" % wb: gc,
                    "c%d_ddirs" % wb: gd})
    np.savez_compressed(os.path.join(OUT, "functions.npz"), **npy(out))


def golden_decoder():
    """TriplanarDecoder.forward + the sampler closure heads (models/generator.py:301-331, 587-681)."""
    B, res, N = 2, 16, 600
    for name, A, use_sdf in (("sdf_attn", 10, True), ("density_rgb", 0, False)):
        planes = synth.make_planes(B, res, seed=1).requires_grad_()
        field = synth.make_field(planes.detach(), seed=3, attention_values=A, use_sdf=use_sdf)
        G = reference_generator(planes, field, 1.4, A=A, use_sdf=use_sdf)
        g = torch.Generator().manual_seed(31)
        pts = ((torch.rand(B, N, 3, generator=g) * 2 - 1) * 1.4 * 1.15)     # some outside the box
        pts[:, :6] = torch.tensor([[1.4, 0, 0], [-1.4, 1.4, 0], [0, 0, 1.4], [1.4, 1.4, 1.4],
                                   [0.0, 0.0, 0.0], [1.3999, -1.4, 0.7]])
        pts = pts.requires_grad_()
        palette = synth.make_palette(B, 10, seed=4).requires_grad_() if A > 0 else None
        dec = G.decoder(planes.view(B, 3, 32, res, res)[:, 0], planes.view(B, 3, 32, res, res)[:, 1],
                        planes.view(B, 3, 32, res, res)[:, 2], (pts / 1.4).view(B, N, 1, 3))
        y = torch.cat((dec["density_or_distance"], dec["features"]), dim=-1)
        gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(32))
        wl = [G.decoder.net[0].weight, G.decoder.net[0].bias, G.decoder.net[2].weight, G.decoder.net[2].bias]
        for p in wl:
            p.requires_grad_()
        dec = G.decoder(planes.view(B, 3, 32, res, res)[:, 0], planes.view(B, 3, 32, res, res)[:, 1],
                        planes.view(B, 3, 32, res, res)[:, 2], (pts / 1.4).view(B, N, 1, 3))
        y = torch.cat((dec["density_or_distance"], dec["features"]), dim=-1)
        gd = torch.autograd.grad((y * gy).sum(), [planes, pts] + wl)
        inputs = {"attention_values": palette} if A > 0 else {}
        mo = G(None, torch.zeros(B, 15 if A > 0 else 14, 512), ["sampler"], inputs)
        so = mo["sampler"](pts.view(B, N, 1, 3), ["sigma", "rgb"] + (["semantics"] if A > 0 else []))
        gsig = torch.randn(so["sigma"].shape, generator=torch.Generator().manual_seed(33))
        grgb = torch.randn(so["rgb"].shape, generator=torch.Generator().manual_seed(34))
        leaves = [planes, pts] + ([palette] if A > 0 else [])
        if use_sdf:
            G.beta.requires_grad_(); G.alpha.requires_grad_()
            mo = G(None, torch.zeros(B, 15 if A > 0 else 14, 512), ["sampler"], inputs)
            so = mo["sampler"](pts.view(B, N, 1, 3), ["sigma", "rgb"] + (["semantics"] if A > 0 else []))
            leaves += [G.beta, G.alpha]
        gs = torch.autograd.grad((so["sigma"] * gsig).sum() + (so["rgb"] * grgb).sum(), leaves)
        out = dict(planes=planes, pts=pts, palette=palette, y=y, gy=gy, **field,
                   dy_planes=gd[0], dy_pts=gd[1], dy_W0=gd[2], dy_b0=gd[3], dy_W1=gd[4], dy_b1=gd[5],
                   sigma=so["sigma"], rgb=so["rgb"], semantics=so.get("semantics"), gsigma=gsig, grgb=grgb,
                   ds_planes=gs[0], ds_pts=gs[1])
        if A > 0:
            out["ds_palette"] = gs[2]
        if use_sdf:
            out["ds_beta"], out["ds_alpha"] = gs[-2], gs[-1]
        np.savez_compressed(os.path.join(OUT, "decoder_%s.npz" % name), **npy(out))


def golden_pose():
    """pose_utils.pose_to_matrix (lib/pose_utils.py:48-75) with F.normalize on the quaternion as the
    inversion loop calls it (run.py:2257-2262), forward and gradients."""
    pu = ref_loader.modules().pose_utils
    g = torch.Generator().manual_seed(41)
    B = 5
    out = {}
    for name, persp, flipped in (("persp_flip", True, True), ("persp", True, False), ("ortho_flip", False, True)):
        B = 5 if persp else 1            # the reference's orthographic branch only works for one image
        z0 = (torch.randn(B, generator=g) * 0.5).requires_grad_() if persp else None
        t2 = (torch.randn(B, 2, generator=g) * 0.3).requires_grad_()
        s = (torch.rand(B, generator=g) + 0.5).requires_grad_()
        q = torch.randn(B, 4, generator=g).requires_grad_()
        mat, focal = pu.pose_to_matrix(z0, t2, s, F.normalize(q, dim=-1), camera_flipped=flipped)
        gm = torch.randn(B, 4, 4, generator=g)
        gf = torch.randn(B, generator=g)
        loss = (mat * gm).sum() + ((focal * gf).sum() if persp else 0)
        leaves = ([z0] if persp else []) + [t2, s, q]
        grads = torch.autograd.grad(loss, leaves)
        d = dict(t2=t2, s=s, q=q, mat=mat, gm=gm, gf=gf)
        if persp:
            d.update(z0=z0, focal=focal, d_z0=grads[0])
        d.update(d_t2=grads[-3], d_s=grads[-2], d_q=grads[-1])
        out.update({"%s_%s" % (name, k): v for k, v in d.items()})
    np.savez_compressed(os.path.join(OUT, "pose.npz"), **npy(out))


def golden_render(only=()):
    """Whole `render` (run.py:176-350; eval_kitti_persp.py:43-230 for the z-buffer variant)."""
    cases = {
        # name: (script, persp, randomize, white, no_cam_grad, scene_range, H, W, center)
        "persp_rand": ("run.py", True, True, False, False, 1.4, 12, 12, False),
        "ortho_det": ("run.py", False, False, False, False, 1.4, 12, 12, False),
        "persp_white_nocam": ("run.py", True, True, True, True, 1.4, 10, 14, False),
        "kitti_zbuf": ("eval_kitti_persp.py", True, True, True, False, 3.0, 12, 12, True),
        # normals + semantics (the evaluation render, eval_kitti_persp.py:538-539)
        # and coords (coord-regressor data synthesis, run.py:1639-1646); forward-only outputs
        "persp_n1": ("run.py", True, True, True, False, 1.4, 12, 12, False),
        "ortho_coords": ("run.py", False, True, False, False, 1.4, 12, 12, False),
    }
    # round 2: rays in multiples of 8 x 16 so that the reference's own gradients pin the pixel-block tile
    # schedule of the warp-specialised kernels (the cases above have H % 8 != 0 and take the linear tile order),
    # on 64 x 64 / 32 x 32 planes (8 x 8 texel blocks), with the parameter sets of BASELINE.json config 3
    cases.update({
        "persp_tiles": ("run.py", True, True, False, False, 1.4, 16, 32, False),
        # ShapeNet (data/loaders.py:120-123, 426-435): scene_range 0.55, white background, one focal, no pose gradient
        "shapenet_tiles": ("run.py", True, True, True, True, 0.55, 16, 16, False),
        # CUB (data/datasets.py:556-567): orthographic with T[3,3] != 1, scene_range 2.0
        "cub_tiles": ("run.py", False, True, False, False, 2.0, 8, 16, False),
    })
    extra = {"persp_tiles": dict(B=1, res=64), "shapenet_tiles": dict(B=2, res=32, camera="shared_focal"),
             "cub_tiles": dict(B=1, res=32, camera="cub")}
    n1 = {"persp_n1": dict(compute_normals=True, compute_semantics=True),
          "ortho_coords": dict(compute_coords=True)}
    for name, (script, persp, randomize, white, no_cam, sr, H, W, use_center) in cases.items():
        if only and name not in only:
            continue
        B, S, res = 2, 64, 16
        if name == "kitti_zbuf":
            B = 1
        ex = extra.get(name, {})
        B, res = ex.get("B", B), ex.get("res", res)
        planes = synth.make_planes(B, res, seed=1).requires_grad_()
        field = synth.make_field(planes.detach(), seed=3)
        palette = synth.make_palette(B, 10, seed=4).requires_grad_()
        if name == "kitti_zbuf":
            cam, _ = synth.make_cameras(B, seed=2, focal_range=(2.0, 2.0), dist_jitter=(2.2, 2.2))
            focal = torch.tensor(3.1)                       # 0-dim, eval_kitti_persp.py:948
            center = torch.tensor([[0.56, 0.47]])
        elif ex.get("camera") == "cub":
            cam, focal, center = synth.make_ortho_cub_cameras(B, seed=7), None, None
        else:
            cam, focal = synth.make_cameras(B, seed=2, perspective=persp)
            center = None
            if ex.get("camera") == "shared_focal":
                cam = cam.clone()
                cam[:, :3, 3] *= sr / 1.4
                focal = focal[:1].expand(B).contiguous()
        cam = cam.detach().requires_grad_()
        if focal is not None:
            focal = focal.detach().requires_grad_()
        G = reference_generator(planes, field, sr)
        render_ref, _ = ref_loader.lift_render(script, scene_range=sr, white_background=white)
        torch.manual_seed(77)
        out = render_ref(G, H, W, cam, focal, center, None, torch.zeros(B, 15, 512), S,
                         randomize=randomize, extra_model_inputs={"attention_values": palette},
                         force_no_cam_grad=no_cam, **n1.get(name, {}))
        torch.manual_seed(77)
        u1 = torch.rand(B, H, W, S) if randomize else None
        u2 = torch.rand(B * H * W, S) if randomize else None
        tgt = torch.rand(B, H, W, 3, generator=torch.Generator().manual_seed(8)) * 2 - 1
        gmask = torch.rand(B, H, W, generator=torch.Generator().manual_seed(9)) - 0.5
        if name in n1:                       # forward-only fixtures
            d = dict(planes=planes, palette=palette, cam=cam, focal=focal, center=center, u_strat=u1,
                     u_pdf=u2, rgb=out[0], depth=out[1], mask=out[2], normals_map=out[3],
                     semantics_map=out[4], **field)
            d["meta"] = np.array([H, W, S, int(persp), int(randomize), int(white), int(no_cam),
                                  int(script != "run.py")], dtype=np.int64)
            d["scene_range"] = np.array(sr, dtype=np.float64)
            d["flags"] = np.array([int(n1[name].get(k, False)) for k in
                                   ("compute_normals", "compute_semantics", "compute_coords")], dtype=np.int64)
            assert float(out[2].mean()) > 0.1, name
            np.savez_compressed(os.path.join(OUT, "render_%s.npz" % name), **npy(d))
            print(name, "mask mean %.3f" % float(out[2].mean()))
            continue
        loss = (out[0] - tgt).abs().mean() * B + (out[2] * gmask).mean()
        wants = [planes, palette] + ([] if no_cam else ([cam] + ([focal] if persp else [])))
        gr = torch.autograd.grad(loss, wants, allow_unused=True)
        d = dict(planes=planes, palette=palette, cam=cam, focal=focal, center=center, u_strat=u1, u_pdf=u2,
                 target=tgt, gmask=gmask, rgb=out[0], depth=out[1], mask=out[2],
                 d_planes=gr[0], d_palette=gr[1], **field)
        d["meta"] = np.array([H, W, S, int(persp), int(randomize), int(white), int(no_cam),
                              int(script != "run.py")], dtype=np.int64)
        d["scene_range"] = np.array(sr, dtype=np.float64)
        if not no_cam:
            d["d_cam"] = gr[2]
            if persp:
                d["d_focal"] = gr[3]
        assert float(out[2].mean()) > 0.1, name
        np.savez_compressed(os.path.join(OUT, "render_%s.npz" % name), **npy(d))
        print(name, "mask mean %.3f" % float(out[2].mean()))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    only = tuple(a for a in sys.argv[1:] if not a.startswith("-"))   # e.g. `make_golden.py persp_n1`
    if not only:
        golden_functions()
        golden_decoder()
    if not only or "pose" in only:
        golden_pose()
    if only != ("pose",):
        golden_render(only)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
