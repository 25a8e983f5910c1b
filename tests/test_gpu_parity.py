"""GPU parity tests: the sm_100a kernels, called through the C ABI (ctypes -> libnfi_b200.so), against
(1) the stored outputs of the unmodified reference (tests/golden/*.npz) and (2) the CPU oracle on
seeded synthetic inputs.  Contract (BASELINE.json north_star): bit-exact ray / sample indices;
rgb, depth, mask and gradients within 1e-4 relative (max-abs-diff / max-abs-ref) in fp32.
"""
import pytest
import torch

import nfi_oracle as O
from golden_io import (NEARFAR_RANGE, RAY_VARIANTS, RENDER_CASES, N1_CASES, eq_bits, load, rel_err,
                       render_case)

pytestmark = pytest.mark.gpu
TOL = 1e-4     # the fp32 tolerance of north_star


def cu(t):
    return None if t is None else t.cuda()


@pytest.fixture(scope="module")
def nfi():
    import nfi_b200
    from nfi_b200 import _lib, decoder, nerf_utils, ops, render, synth
    _lib.lib()     # fails loudly if the CUDA library is missing
    import types
    return types.SimpleNamespace(ops=ops, nu=nerf_utils, dec=decoder, render=render, synth=synth)


@pytest.fixture(scope="module")
def fn():
    return load("functions")


# ---------------------------------------------------------------- A1
@pytest.mark.parametrize("variant", sorted(RAY_VARIANTS))
def test_rays_bitexact(nfi, fn, variant):
    cam_key, persp, bb, c = RAY_VARIANTS[variant]
    H, W = [int(v) for v in fn["HW"]]
    o, d = nfi.nu.get_ray_bundle(H, W, cu(fn["focal"]) if persp else None, cu(fn[cam_key]),
                                 cu(fn["bbox"]) if bb else None, cu(fn["center"]) if c else None)
    assert eq_bits(o.cpu(), fn["rays_o_" + variant].expand_as(fn["rays_d_" + variant]))
    assert eq_bits(d.cpu(), fn["rays_d_" + variant])


def test_rays_scalar_focal(nfi, fn):
    H, W = [int(v) for v in fn["HW"]]
    _, d = nfi.nu.get_ray_bundle(H, W, cu(fn["focal"][0]), cu(fn["cam_persp"][:1]), None,
                                 cu(fn["center"][:1]))
    assert eq_bits(d.cpu(), fn["rays_d_persp_scalar_focal"])


@pytest.mark.parametrize("kind", ["persp", "ortho", "cub"])
def test_rays_backward(nfi, fn, kind):
    cam_key, persp, _, _ = RAY_VARIANTS[kind]
    H, W = [int(v) for v in fn["HW"]]
    cam = fn[cam_key].clone().requires_grad_()
    focal = fn["focal"].clone().requires_grad_() if persp else None
    o, d = O.pixel_rays(H, W, cam, focal, fn["center"] if persp else None)
    d = O.unit_directions(d)
    g = torch.Generator().manual_seed(5)
    go, gd = torch.randn(o.expand_as(d).shape, generator=g), torch.randn(d.shape, generator=g)
    ref = torch.autograd.grad((o.expand_as(d) * go).sum() + (d * gd).sum(), [cam] + ([focal] if persp else []))
    cam2 = cu(fn[cam_key]).requires_grad_()
    focal2 = cu(fn["focal"]).requires_grad_() if persp else None
    o2, d2 = nfi.ops.ray_bundle(H, W, focal2, cam2, None, cu(fn["center"]) if persp else None, normalize=True)
    got = torch.autograd.grad((o2 * cu(go)).sum() + (d2 * cu(gd)).sum(), [cam2] + ([focal2] if persp else []))
    for a, b in zip(got, ref):
        assert rel_err(a.cpu(), b) < TOL


# ---------------------------------------------------------------- A2, A3
@pytest.mark.parametrize("variant", sorted(NEARFAR_RANGE))
def test_near_far_bitexact(nfi, fn, variant):
    cam_key, persp, _, _ = RAY_VARIANTS[variant]
    H, W = [int(v) for v in fn["HW"]]
    o, d = nfi.ops.ray_bundle(H, W, cu(fn["focal"]) if persp else None, cu(fn[cam_key]), None, None,
                              normalize=True)
    assert eq_bits(d.cpu(), fn["unit_d_" + variant])
    n, f = nfi.nu.compute_near_far_planes(o, d, NEARFAR_RANGE[variant])
    assert eq_bits(n.cpu(), fn["near_" + variant]) and eq_bits(f.cpu(), fn["far_" + variant])


def test_near_far_all_miss_raises(nfi):
    # same rays as tests/test_oracle_vs_reference.py::test_near_far_axis_aligned_and_all_miss:
    # axis-aligned directions (exact zeros -> +-inf reciprocals), one ray outside the box
    o = torch.tensor([[0.0, 0.0, 3.0], [0.2, -0.3, 3.0], [5.0, 0.0, 3.0], [1.4, 0.0, 3.0]])
    d = torch.tensor([[0.0, 0.0, -1.0], [0.0, -0.0, -1.0], [0.0, 0.0, -1.0], [0.0, 0.0, -1.0]])
    n_ref, f_ref = O.box_near_far(o, d, 1.4)
    n, f = nfi.nu.compute_near_far_planes(o.cuda(), d.cuda(), 1.4)
    assert eq_bits(n.cpu(), n_ref) and eq_bits(f.cpu(), f_ref)
    with pytest.raises(RuntimeError):
        O.box_near_far(o[2:3], d[2:3], 1.4)
    with pytest.raises(RuntimeError):       # the reference raises here too (nerf_utils.py:260)
        nfi.nu.compute_near_far_planes(o[2:3].cuda(), d[2:3].cuda(), 1.4)


def test_stratified_bitexact(nfi, fn):
    n, f = cu(fn["near_persp"]), cu(fn["far_persp"])
    assert eq_bits(nfi.ops.stratified(n, f, 64, False).cpu(), fn["strat_z_det"])
    z = nfi.ops.stratified(n, f, 64, True, cu(fn["strat_u"]))
    assert eq_bits(z.cpu(), fn["strat_z_rand"])
    H, W = [int(v) for v in fn["HW"]]
    o, d = nfi.ops.ray_bundle(H, W, cu(fn["focal"]), cu(fn["cam_persp"]), None, None, normalize=True)
    assert eq_bits(nfi.ops.points_on_rays(o, d, z).cpu(), fn["strat_pts_rand"])


def test_stratified_philox_statistics(nfi, fn):
    """perf mode (no uniforms passed): in-kernel Philox jitter is uniform in [0, 1) per stratum."""
    n, f = cu(fn["near_persp"]), cu(fn["far_persp"])
    z0 = nfi.ops.stratified(n, f, 64, False)
    z = nfi.ops.stratified(n, f, 64, True, None, seed=1234, offset=7)
    t = ((z - z0) / ((f - n) / 64).unsqueeze(-1)).flatten()
    assert float(t.min()) >= 0.0 and float(t.max()) < 1.0 + 1e-4
    assert abs(float(t.mean()) - 0.5) < 5e-3 and abs(float(t.var()) - 1 / 12) < 5e-3
    z2 = nfi.ops.stratified(n, f, 64, True, None, seed=1234, offset=7)
    assert eq_bits(z.cpu(), z2.cpu())


# ---------------------------------------------------------------- A7, A8
def test_weights_smoothing_pdf(nfi, fn):
    """Float values within tolerance (CUDA expf and the CPU's differ by an ulp); the sample INDICES
    and depths of sample_pdf are bit-exact for identical inputs (ATen's reduction order is
    reproduced in the kernel)."""
    sg, zz, dd = cu(fn["w_sigma"]), cu(fn["w_z"]), cu(fn["w_dirs"])
    w = nfi.nu.render_volume_density_weights_only(sg, None, dd, zz)
    assert rel_err(w.cpu(), fn["w_weights"]) < 1e-6
    ws = nfi.ops.weights_only(sg, dd, zz, smooth=True)
    assert rel_err(ws.cpu(), fn["w_smooth"]) < 1e-6
    mid = 0.5 * (zz[..., 1:] + zz[..., :-1])
    mid = mid.reshape(-1, mid.shape[-1])
    wsel = cu(fn["w_smooth"]).reshape(-1, ws.shape[-1])[..., 1:-1].contiguous()     # identical inputs
    zr, inds = nfi.ops.sample_pdf(mid, wsel, 64, cu(fn["pdf_u"]), return_indices=True)
    assert eq_bits(zr.cpu(), fn["pdf_z_rand"])
    _, ref_inds, _, _ = O.inverse_cdf_samples(mid.cpu(), wsel.cpu(), 64, fn["pdf_u"], return_indices=True)
    assert bool((inds.cpu().long() == ref_inds).all())          # bit-exact sample indices
    assert eq_bits(nfi.ops.sample_pdf(mid, wsel, 64, None).cpu(), fn["pdf_z_det"])
    # the fused importance kernel (weights -> smoothing -> pdf in one pass): its weights carry the
    # expf ulp differences, so a sample may flip to the neighbouring bin only where u is within
    # rounding of a CDF edge -- count those
    zf = nfi.ops.importance(sg, dd, zz, cu(fn["pdf_u"])).reshape(-1, 64).cpu()
    diff = (zf - fn["pdf_z_rand"]).abs()
    scale = float(fn["pdf_z_rand"].abs().max())
    assert float((diff > 1e-5 * scale).float().mean()) < 1e-3
    assert float(diff.max()) < 0.1 * scale


def test_cumprod_exclusive(nfi):
    x = torch.rand(37, 200, generator=torch.Generator().manual_seed(3)) * 0.2 + 0.85
    ref = torch.cumprod(x.double(), -1).roll(1, -1)
    ref[..., 0] = 1
    assert rel_err(nfi.nu.cumprod_exclusive(x.cuda()).cpu(), ref.float()) < 1e-6


# ---------------------------------------------------------------- A10
@pytest.mark.parametrize("white", [0, 1])
def test_composite(nfi, fn, white):
    sig = cu(fn["c_sigma"]).requires_grad_()
    col = cu(fn["c_rgb"]).requires_grad_()
    dirs = cu(fn["c_dirs"]).requires_grad_()
    sem = cu(fn["c_sem"]).requires_grad_()
    out = nfi.nu.render_volume_density(sig, col, None, dirs, cu(fn["c_z"]), cu(fn["c_nrm"]), sem,
                                       white_background=bool(white))
    for k, v in zip(("rgb", "depth", "mask", "nrm", "sem"), out):
        assert rel_err(v.cpu(), fn["c%d_%s" % (white, k)]) < 1e-5, k
    gs, gc, gd = torch.autograd.grad((out[0] * cu(fn["c_grad_rgb"])).sum() + (out[2] * cu(fn["c_grad_mask"])).sum(),
                                     (sig, col, dirs))
    assert rel_err(gs.cpu(), fn["c%d_dsigma" % white]) < TOL
    assert rel_err(gc.cpu(), fn["c%d_drgb" % white]) < TOL
    assert rel_err(gd.cpu(), fn["c%d_ddirs" % white]) < TOL


def test_composite_semantics_backward(nfi, fn):
    g = torch.Generator().manual_seed(11)
    sig = fn["c_sigma"].clone().requires_grad_()
    col = fn["c_rgb"].clone().requires_grad_()
    sem = fn["c_sem"].clone().requires_grad_()
    dn = O.vec3_norm(fn["c_dirs"])
    out = O.composite(sig, col, fn["c_z"], dn, False, None, sem)
    gsem = torch.randn(out[4].shape, generator=g)
    ref = torch.autograd.grad((out[4] * gsem).sum() + out[0].sum(), (sig, sem))
    sig2, col2, sem2 = cu(fn["c_sigma"]).requires_grad_(), cu(fn["c_rgb"]).requires_grad_(), cu(fn["c_sem"]).requires_grad_()
    out2 = nfi.ops.composite(sig2, col2, cu(fn["c_dirs"]), cu(fn["c_z"]), None, sem2, False)
    got = torch.autograd.grad((out2[4] * cu(gsem)).sum() + out2[0].sum(), (sig2, sem2))
    for a, b in zip(got, ref):
        assert rel_err(a.cpu(), b) < TOL


# ---------------------------------------------------------------- A4-A6
def _field(nfi, g, A, use_sdf, dev="cuda", grad=False):
    mk = (lambda t: t.clone().to(dev).requires_grad_()) if grad else (lambda t: t.to(dev))
    return nfi.dec.Field(mk(g["W0"]), mk(g["b0"]), mk(g["W1"]), mk(g["b1"]),
                         mk(g["beta"]) if use_sdf else None, mk(g["alpha"]) if use_sdf else None,
                         1.4, A, use_sdf)


@pytest.mark.parametrize("name,A,use_sdf", [("sdf_attn", 10, True), ("density_rgb", 0, False)])
def test_decoder_and_heads(nfi, name, A, use_sdf):
    g = load("decoder_" + name)
    B, res = g["planes"].shape[0], g["planes"].shape[-1]
    planes = cu(g["planes"]).requires_grad_()
    pts = cu(g["pts"]).requires_grad_()
    fld = _field(nfi, g, A, use_sdf)
    # raw decoder (TriplanarDecoder.forward): coords are in box units
    pcl = nfi.ops.planes_to_channels_last(planes)
    cfg_raw = nfi.ops.FieldConfig(A, False, 1.0)
    y, _ = nfi.ops.field_on_points(pcl, None, fld.W0, fld.b0, fld.W1, fld.b1, None, None, pts / 1.4,
                                   cfg_raw, mode=1)
    assert rel_err(y.cpu(), g["y"]) < TOL
    gd = torch.autograd.grad((y * cu(g["gy"])).sum(), [planes, pts])
    assert rel_err(gd[0].cpu(), g["dy_planes"]) < TOL
    assert rel_err(gd[1].cpu(), g["dy_pts"]) < TOL
    # decoder weight gradients (generator training, run.py:980-986)
    wl = [t.clone().requires_grad_() for t in (fld.W0, fld.b0, fld.W1, fld.b1)]
    yw, _ = nfi.ops.field_on_points(pcl, None, *wl, None, None, pts / 1.4, cfg_raw, mode=1)
    gw = torch.autograd.grad((yw * cu(g["gy"])).sum(), [planes] + wl)
    assert rel_err(gw[0].cpu(), g["dy_planes"]) < TOL
    for a, k in zip(gw[1:], ("dy_W0", "dy_b0", "dy_W1", "dy_b1")):
        assert rel_err(a.cpu(), g[k]) < TOL, k
    # heads
    if use_sdf:     # beta / alpha are trainable in G training (generator.py:397-399)
        fld.beta, fld.alpha = fld.beta.clone().requires_grad_(), fld.alpha.clone().requires_grad_()
    pal = cu(g["palette"]).requires_grad_() if A > 0 else None
    o4, probs = nfi.ops.field_on_points(pcl, pal, *fld.tensors(), pts, fld.cfg, mode=0, want_probs=True)
    sigma, rgb = o4[..., 0], o4[..., 1:]
    assert rel_err(sigma.cpu(), g["sigma"]) < TOL and rel_err(rgb.cpu(), g["rgb"]) < TOL
    if A > 0:
        assert rel_err(probs.cpu(), g["semantics"]) < TOL
    leaves = [planes, pts] + ([pal] if A > 0 else []) + ([fld.beta, fld.alpha] if use_sdf else [])
    keys = ["ds_planes", "ds_pts"] + (["ds_palette"] if A > 0 else []) + (["ds_beta", "ds_alpha"] if use_sdf else [])
    gs = torch.autograd.grad((sigma * cu(g["gsigma"])).sum() + (rgb * cu(g["grgb"])).sum(), leaves)
    # ds_alpha = -(1/alpha) * sum(gsigma * sigma) over random-sign gsigma is an ill-conditioned fp32 sum
    # (sum|terms| / |sum| = 1.3e3 on this fixture), so the fp32 golden itself is only good to a few
    # eps * sum|terms|; that floor is added to the contract tolerance for the two scalar reductions.
    floor = 8 * 1.1920929e-07 * float((torch.as_tensor(g["sigma"]) * torch.as_tensor(g["gsigma"])).abs().sum())
    for a, k in zip(gs, keys):
        if k in ("ds_alpha", "ds_beta"):
            ref = torch.as_tensor(g[k]).double()
            assert float((a.cpu().double() - ref).abs().max()) < TOL * float(ref.abs().max()) + floor, k
        else:
            assert rel_err(a.cpu(), g[k]) < TOL, k


def test_sampler_dropin(nfi):
    g = load("decoder_sdf_attn")
    fld = _field(nfi, g, 10, True)
    smp = nfi.dec.make_sampler(cu(g["planes"]), cu(g["palette"]), fld)
    out = smp(cu(g["pts"]).unsqueeze(2), ["sigma", "rgb", "semantics", "sdf_distance"])
    assert rel_err(out["sigma"].cpu(), g["sigma"]) < TOL and rel_err(out["rgb"].cpu(), g["rgb"]) < TOL
    assert rel_err(out["semantics"].cpu(), g["semantics"]) < TOL
    assert rel_err(out["sdf_distance"].cpu(), g["y"][..., :1]) < TOL


# ---------------------------------------------------------------- A11: the whole render
def _run_render(nfi, g, c, planes, pal, cam, focal):
    fld = nfi.dec.Field(cu(g["W0"]), cu(g["b0"]), cu(g["W1"]), cu(g["b1"]), cu(g["beta"]), cu(g["alpha"]),
                        c["scene_range"], 10, True)
    return nfi.render.render_planes(
        planes, pal, fld, c["H"], c["W"], cam, focal, cu(g.get("center")), None, c["S"],
        randomize=c["randomize"], white_background=c["white"], fine_sampling=True,
        force_no_cam_grad=c["no_cam"], zbuffer_depth=c["zbuffer"], u_strat=cu(g.get("u_strat")),
        u_pdf=cu(g.get("u_pdf")), return_aux=True)


@pytest.mark.parametrize("name", RENDER_CASES)
def test_render_golden(nfi, name):
    g = render_case(name)
    c = g["cfg"]
    B = g["planes"].shape[0]
    planes = cu(g["planes"]).requires_grad_()
    pal = cu(g["palette"]).requires_grad_()
    cam = cu(g["cam"]).requires_grad_()
    focal = cu(g["focal"]).requires_grad_() if c["persp"] else None
    rgb, depth, mask, aux = _run_render(nfi, g, c, planes, pal, cam, focal)
    assert rel_err(rgb.cpu(), g["rgb"]) < TOL and rel_err(depth.cpu(), g["depth"]) < TOL
    assert rel_err(mask.cpu(), g["mask"]) < TOL
    loss = (rgb - cu(g["target"])).abs().mean() * B + (mask * cu(g["gmask"])).mean()
    wants = [planes, pal] + ([] if c["no_cam"] else [cam] + ([focal] if c["persp"] else []))
    keys = ["d_planes", "d_palette"] + ([] if c["no_cam"] else ["d_cam"] + (["d_focal"] if c["persp"] else []))
    for a, k in zip(torch.autograd.grad(loss, wants), keys):
        assert rel_err(a.cpu(), g[k]) < TOL, k


# ---------------------------------------------------------------- decoder MLP on tcgen05
@pytest.mark.parametrize("n", [128, 1000, 128 * 300 + 17])
def test_decoder_mlp_tensor_cores(nfi, n):
    g = torch.Generator().manual_seed(21)
    feat = torch.randn(n, 32, generator=g)
    W0, b0 = torch.randn(64, 32, generator=g), torch.randn(64, generator=g) * 0.1
    W1, b1 = torch.randn(11, 64, generator=g), torch.randn(11, generator=g) * 0.1
    hp_ref = torch.nn.functional.linear(feat.double(), W0.double() / 32 ** 0.5, b0.double())
    y_ref = torch.nn.functional.linear(torch.nn.functional.softplus(hp_ref), W1.double() / 8.0, b1.double())
    y, hp = nfi.ops.decoder_mlp(cu(feat), cu(W0), cu(b0), cu(W1), cu(b1), want_hidden=True)
    assert rel_err(hp.cpu(), hp_ref) < 5e-6
    assert rel_err(y.cpu(), y_ref) < 5e-6


# ---------------------------------------------------------------- A11/A12 against the live oracle
def _oracle_vs_cuda(nfi, B, H, W, S, res, perspective, train, white, no_cam, seed=0, oracle_dev="cpu",
                    plane_dtype=torch.float32, oracle_bf16_planes=False, loss_kind="l1", feed_z_fine=False):
    """One render forward + backward through the C ABI and through the CPU oracle on the same seeded
    inputs; returns {name: relative error}.  train=True is the generator-training gradient set
    (planes, attention values, W0, b0, W1, b1, beta, alpha; poses are data)."""
    import nfi_oracle as O
    from nfi_b200 import synth
    planes = synth.make_planes(B, res=res, seed=1 + seed)
    fd = synth.make_field(planes, seed=3 + seed)
    pal = synth.make_palette(B, 10, seed=4 + seed)
    cam, focal = synth.make_cameras(B, seed=2 + seed, perspective=perspective)
    g = torch.Generator().manual_seed(99 + seed)
    u_s = torch.rand(B, H, W, S, generator=g)
    u_p = torch.rand(B * H * W, S, generator=g)
    target = torch.rand(B, H, W, 3, generator=g) * 2 - 1
    wkeys = ("W0", "b0", "W1", "b1", "beta", "alpha")

    def run(dev, use_oracle, z_fine=None):
        lv = {"planes": planes, "palette": pal}
        if not train and not no_cam:
            lv["cam"] = cam
            if perspective:
                lv["focal"] = focal
        if train:
            lv.update({k: fd[k] for k in wkeys})
        if use_oracle and oracle_bf16_planes:      # the oracle sees what a bf16 plane copy holds
            lv["planes"] = planes.to(torch.bfloat16).float()
        lv = {k: v.clone().to(dev).requires_grad_() for k, v in lv.items()}
        w = {k: lv.get(k, fd[k].to(dev)) for k in wkeys}
        c = lv.get("cam", cam.to(dev))
        f = (lv.get("focal", focal.to(dev)) if perspective else None)
        if use_oracle:
            fp = O.FieldParams(w["W0"], w["b0"], w["W1"], w["b1"], 1.4, True, 10, w["beta"], w["alpha"])
            out = O.render(lv["planes"].view(B, 3, 32, res, res), lv["palette"], fp, H, W, c, f, None, None, S,
                           white, u_s.to(dev), u_p.to(dev), True, True, no_cam, return_aux=True)
        else:
            fld = nfi.dec.Field(*[w[k] for k in wkeys], 1.4, 10, True)
            out = nfi.render.render_planes(lv["planes"], lv["palette"], fld, H, W, c, f, None, None, S,
                                           randomize=True, white_background=white, force_no_cam_grad=no_cam,
                                           u_strat=u_s.to(dev), u_pdf=u_p.to(dev), return_aux=True,
                                           plane_dtype=plane_dtype, z_fine=z_fine)
        rgb, depth, mask = out[:3]
        diff = rgb - target.to(dev)
        loss = (diff.abs() if loss_kind == "l1" else diff * diff).mean() * B + 0.1 * mask.mean()
        grads = torch.autograd.grad(loss, list(lv.values()))
        res_ = {"rgb": rgb, "depth": depth, "mask": mask}
        res_.update({"d_" + k: gr for k, gr in zip(lv, grads)})
        return res_, out[3]["z_fine"].detach().cpu().reshape(-1)

    ref, zf_ref = run(oracle_dev, True)
    got, zf_got = run("cuda", False, zf_ref.view(B, H, W, S).cuda() if feed_z_fine else None)
    assert float(ref["mask"].detach().mean()) > 0.05, "vacuous fixture"
    errs = {k: rel_err(got[k].detach().cpu(), ref[k].detach().cpu()) for k in ref}
    # importance samples that moved by a whole bin between the two summation orders
    flips = int(((zf_got - zf_ref).abs() > 1e-4 * zf_ref.abs().clamp_min(1.0)).sum())
    return errs, flips, zf_ref.numel()


# Pose gradients (d_cam, d_focal) are sums over every sample of a piecewise-constant quantity: the
# bilinear coordinate gradient jumps at texel boundaries.  Of the ~1e5 fine samples of an image a few
# land within an ulp or two of a boundary (their depth comes out of the inverse CDF, which differs
# by ulps between the CPU and the GPU exp), floor() puts them in different cells, and that one sample
# moves the pose gradient of its ray by O(1) (tools/dev/cam_probe2.py shows the rays).  The committed
# reference fixtures (test_render_golden) hold d_cam / d_focal to 1e-4; here, on fresh seeds, they get
# the looser bound and everything else the contract tolerance.
POSE_TOL = 5e-3


def _check(errs, flips, n):
    assert flips <= max(2, n // 100000), "too many importance-sample mismatches: %d of %d" % (flips, n)
    for k, v in errs.items():
        tol = POSE_TOL if k in ("d_cam", "d_focal") else TOL
        assert v < tol, (k, v, flips)


@pytest.mark.parametrize("B,HW,seed", [(1, 32, 0), (3, 32, 1), (1, 64, 2), (2, 48, 3)])
def test_pose_gradients_meet_the_contract_on_identical_fine_samples(nfi, B, HW, seed):
    """POSE_TOL above is an argument; this is the measurement behind it.  With the ORACLE's fine depths fed
    to the CUDA path (render_planes(z_fine=...)) the only difference left is the differentiable path itself,
    and d_cam / d_focal then meet the 1e-4 contract like everything else -- so the looser bound of the
    fresh-seed tests is provably importance-bin flips (few samples whose depth comes out of the inverse CDF a
    whole bin away), not the gradient kernels."""
    errs, flips, n = _oracle_vs_cuda(nfi, B, HW, HW, 64, 64, True, False, False, False, seed=seed, feed_z_fine=True)
    assert flips == 0
    for k, v in errs.items():
        assert v < TOL, (k, v)


@pytest.mark.parametrize("B,HW,perspective,white,no_cam", [
    (1, 32, True, False, False), (3, 32, True, True, False), (2, 32, False, False, False),
    (2, 16, True, False, True), (1, 64, True, False, False)])
def test_render_vs_oracle_sizes(nfi, B, HW, perspective, white, no_cam):
    _check(*_oracle_vs_cuda(nfi, B, HW, HW, 64, 64, perspective, False, white, no_cam))


def test_render_training_gradients_vs_oracle(nfi):
    """Generator-training gradient set: decoder weights, beta, alpha, planes, attention values."""
    errs, flips, n = _oracle_vs_cuda(nfi, 2, 32, 32, 64, 64, True, True, False, False, seed=5)
    _check(errs, flips, n)
    assert {"d_W0", "d_b0", "d_W1", "d_b1", "d_beta", "d_alpha"} <= set(errs)


# ---------------------------------------------------------------- bf16 mode (north_star: 1e-2)
@pytest.mark.parametrize("B,HW,perspective", [(2, 32, True), (1, 64, True), (2, 32, False)])
def test_render_bf16_planes_vs_oracle_on_rounded_planes(nfi, B, HW, perspective):
    """plane_dtype=torch.bfloat16 changes ONE thing: the gather reads a bfloat16 copy of the planes.  Against
    the oracle fed with the same rounded planes everything must therefore still hold to the fp32 tolerance
    (images, importance samples, and every gradient -- the plane gradient stays fp32)."""
    _check(*_oracle_vs_cuda(nfi, B, HW, HW, 64, 64, perspective, False, False, False, seed=7,
                            plane_dtype=torch.bfloat16, oracle_bf16_planes=True))


BF16_TOL = 1e-2          # the bf16-mode tolerance of north_star
BF16_POSE_TOL = 5e-2     # see below


def test_render_bf16_mode_contract(nfi):
    """bf16 mode against the fp32 reference path on the ORIGINAL planes: rgb, depth, mask, plane and
    attention-value gradients within 1e-2 relative (smooth loss: an L1 loss's sign() would turn a 1e-3
    change of rgb into an O(1) change of single gradient entries in either implementation).  Measured:
    images 6e-4, d_planes 3e-3.  The pose gradients (d_cam, d_focal) are sums of texel DIFFERENCES
    (dF/dx = v1 - v0 ...): rounding smooth planes to 8 mantissa bits leaves those differences with a few
    percent of relative error whatever the arithmetic after the load does (measured 3e-2), so they get the
    looser bound -- a property of storing planes in bf16, stated rather than hidden."""
    errs, _, _ = _oracle_vs_cuda(nfi, 2, 32, 32, 64, 64, True, False, False, False, seed=7,
                                 plane_dtype=torch.bfloat16, loss_kind="l2")
    for k, v in errs.items():
        assert v < (BF16_POSE_TOL if k in ("d_cam", "d_focal") else BF16_TOL), (k, v, errs)


def test_bf16_planes_with_weight_gradients(nfi):
    """Round 2: the generator-training gradient set runs on the warp-specialised kernel (nfi_field_bwd_train), which
    has the bf16 gather -- a bf16 plane copy and decoder weight gradients together, held to the fp32 tolerance
    against the oracle fed with the same rounded planes (round 1 rejected the combination)."""
    errs, flips, n = _oracle_vs_cuda(nfi, 2, 16, 32, 64, 64, True, True, False, False, seed=7,
                                     plane_dtype=torch.bfloat16, oracle_bf16_planes=True)
    assert {"d_W0", "d_b0", "d_W1", "d_b1", "d_beta", "d_alpha"} <= set(errs)
    _check(errs, flips, n)


# ---------------------------------------------------------------- BASELINE.json sizes
def test_render_config2_slice_vs_eager_cuda(nfi):
    """Two images of config 2 at its real size (128x128 rays, 64+64 samples, 256x256 planes, perspective,
    pose gradients) against the oracle run with the reference's own ATen ops on the same GPU -- the
    reference's eager CUDA path (SURVEY.md 8(c) "oracle device"), which finishes in seconds where the
    CPU oracle would not."""
    errs, flips, n = _oracle_vs_cuda(nfi, 2, 128, 128, 64, 256, True, False, False, False, seed=11,
                                     oracle_dev="cuda")
    _check(errs, flips, n)


def test_render_config2_full_vs_eager_cuda(nfi):
    """BASELINE.json config 2 at its full size -- 16 images, 128x128 rays, 64+64 samples, 256x256 planes,
    gradients to planes / attention values / pose / focal -- against the eager-CUDA oracle (the reference's
    own ATen ops on the same GPU; ~25 GB of activations, which a 180 GB B200 holds).

    Images, mask, depth and the small gradients are held to the contract tolerance.  The plane gradient
    is compared texel by texel: at this size (33.5 M samples) any two fp32 implementations of the path --
    CPU oracle, eager CUDA, these kernels -- differ in a handful of ISOLATED single-sample events per
    image (a 2 x 2 x 3-plane footprint whose contribution lands differently; tools/dev/sample_event_probe.py
    shows one where the kernels agree with the CPU oracle to all printed digits and eager CUDA is the odd
    one out, and one the other way round), each worth ~1e-2 of the gradient's max-norm because a texel
    only collects a few samples.  So: at most 0.1 % of the texels of an image may be off by more than the
    contract tolerance, and the gradient as a whole must agree to 5e-3 in the Euclidean norm (one event
    is worth ~1e-3 of it)."""
    import nfi_oracle as O
    from nfi_b200 import synth
    B, H, W, S, res = 16, 128, 128, 64, 256
    planes = synth.make_planes(B, res=res, seed=1)
    fd = synth.make_field(planes, seed=3)
    pal = synth.make_palette(B, 10, seed=4)
    cam, focal = synth.make_cameras(B, seed=2, perspective=True)
    g = torch.Generator().manual_seed(99)
    u_s, u_p = torch.rand(B, H, W, S, generator=g).cuda(), torch.rand(B * H * W, S, generator=g).cuda()
    target = (torch.rand(B, H, W, 3, generator=g) * 2 - 1).cuda()
    wk = ("W0", "b0", "W1", "b1", "beta", "alpha")
    w = {k: fd[k].cuda() for k in wk}

    def run(ours):
        lv = [t.clone().cuda().requires_grad_() for t in (planes, pal, cam, focal)]
        if ours:
            fld = nfi.dec.Field(*[w[k] for k in wk], 1.4, 10, True)
            rgb, depth, mask = nfi.render.render_planes(lv[0], lv[1], fld, H, W, lv[2], lv[3], None, None, S,
                                                        randomize=True, u_strat=u_s, u_pdf=u_p)
        else:
            fp = O.FieldParams(w["W0"], w["b0"], w["W1"], w["b1"], 1.4, True, 10, w["beta"], w["alpha"])
            rgb, depth, mask = O.render(lv[0].view(B, 3, 32, res, res), lv[1], fp, H, W, lv[2], lv[3], None, None,
                                        S, False, u_s, u_p)
        loss = ((rgb - target) ** 2).sum() / (H * W * 3) + 0.1 * mask.mean()     # smooth: no sign() events
        return (rgb, depth, mask), torch.autograd.grad(loss, lv)

    (out_r, g_r), (out_o, g_o) = run(False), run(True)
    assert float(out_r[2].mean()) > 0.05, "vacuous fixture"
    for a, b_, name in zip(out_o, out_r, ("rgb", "depth", "mask")):
        assert rel_err(a.detach().cpu(), b_.detach().cpu()) < TOL, name
    # each attention-value gradient is an fp32 sum over the ~2 M samples of its image in both implementations
    # (theirs a bmm reduction, ours per-thread partial sums + one atomic per image change): 5e-4 at this length
    assert rel_err(g_o[1].cpu(), g_r[1].cpu()) < 5e-4, ("d_palette", rel_err(g_o[1].cpu(), g_r[1].cpu()))
    assert rel_err(g_o[2].cpu(), g_r[2].cpu()) < POSE_TOL and rel_err(g_o[3].cpu(), g_r[3].cpu()) < POSE_TOL
    gp_o, gp_r = g_o[0].double().cpu(), g_r[0].double().cpu()
    for b in range(B):
        den = gp_r[b].abs().max()
        per_texel = (gp_o[b] - gp_r[b]).abs().view(3, 32, res, res).amax(dim=1)
        frac = float((per_texel > TOL * den).double().mean())
        assert frac < 1e-3, "image %d: %.3g of the texels off by more than %g of the max-norm" % (b, frac, TOL)
        l2 = float((gp_o[b] - gp_r[b]).norm() / gp_r[b].norm())
        assert l2 < 5e-3, "image %d: relative L2 error %.3g" % (b, l2)


def test_render_config2_full_batch_sharding_property(nfi):
    """Config 2 at full size (B=16, 128x128, 64+64, 256x256 planes): a size-independent property instead
    of an oracle run.  Images are independent, so rendering the batch in one call and in
    two shards of 8 with the same uniforms must give the same rgb / depth / mask bit for bit (the
    miss-ray near/far fill differs per call but only touches samples outside the box, whose density is
    masked to exactly 0) and the same gradients up to the order of the fp32 atomics; and the plane
    gradient is linear in the upstream loss scale."""
    from nfi_b200 import synth
    B, H, W, S, res = 16, 128, 128, 64, 256
    planes = synth.make_planes(B, res=res, seed=1)
    fd = synth.make_field(planes, seed=3)
    pal = synth.make_palette(B, 10, seed=4)
    cam, focal = synth.make_cameras(B, seed=2, perspective=True)
    g = torch.Generator().manual_seed(5)
    target = (torch.rand(B, H, W, 3, generator=g) * 2 - 1).cuda()
    u_s = torch.rand(B, H, W, S, generator=g).cuda()
    u_p = torch.rand(B * H * W, S, generator=g).cuda()
    fld = nfi.dec.Field(*[fd[k].cuda() for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], 1.4, 10, True)

    def run(lo, hi, scale=1.0):
        lv = [t[lo:hi].clone().cuda().requires_grad_() for t in (planes, pal, cam, focal)]
        rgb, depth, mask = nfi.render.render_planes(
            lv[0], lv[1], fld, H, W, lv[2], lv[3], None, None, S, randomize=True, white_background=False,
            u_strat=u_s[lo:hi], u_pdf=u_p[lo * H * W:hi * H * W])
        loss = scale * (rgb - target[lo:hi]).abs().sum() / (H * W * 3)
        return (rgb, depth, mask), torch.autograd.grad(loss, lv)

    (full, gfull) = run(0, B)
    parts = [run(0, 8), run(8, 16)]
    assert float(full[2].mean()) > 0.05, "vacuous fixture"
    for i, name in enumerate(("rgb", "depth", "mask")):
        cat = torch.cat([p[0][i] for p in parts])
        assert torch.equal(full[i], cat), name
    for i, name in enumerate(("d_planes", "d_palette", "d_cam", "d_focal")):
        cat = torch.cat([p[1][i] for p in parts])
        assert rel_err(cat.cpu(), gfull[i].cpu()) < 1e-5, name
    _, g2 = run(0, B, scale=2.0)
    assert rel_err(g2[0].cpu(), 2 * gfull[0].cpu()) < 1e-5


# ---------------------------------------------------------------- SURVEY.md 8(f) N1
NORMALS_TOL = 5e-4    # unit vectors g/|g|: where |g| is small the 1e-5 relative error of g is amplified


@pytest.mark.parametrize("name", N1_CASES)
def test_render_normals_semantics_coords_golden(nfi, name):
    """compute_normals / compute_semantics / compute_coords (run.py:186-188) against the unmodified
    reference's render outputs."""
    g = render_case(name)
    c = g["cfg"]
    wn, ws, wc = [bool(v) for v in g["flags"]]
    fld = nfi.dec.Field(cu(g["W0"]), cu(g["b0"]), cu(g["W1"]), cu(g["b1"]), cu(g["beta"]), cu(g["alpha"]),
                        c["scene_range"], 10, True)
    focal = cu(g["focal"]) if c["persp"] else None
    rgb, depth, mask, aux = nfi.render.render_planes(
        cu(g["planes"]), cu(g["palette"]), fld, c["H"], c["W"], cu(g["cam"]), focal, cu(g.get("center")), None,
        c["S"], randomize=c["randomize"], white_background=c["white"], u_strat=cu(g.get("u_strat")),
        u_pdf=cu(g.get("u_pdf")), compute_normals=wn, compute_semantics=ws, compute_coords=wc)
    assert rel_err(rgb.cpu(), g["rgb"]) < TOL and rel_err(depth.cpu(), g["depth"]) < TOL
    assert rel_err(mask.cpu(), g["mask"]) < TOL
    if wn:
        assert rel_err(aux["normals"].cpu(), g["normals_map"]) < NORMALS_TOL
    else:
        assert aux["normals"] is None
    assert aux["semantics"].shape == g["semantics_map"].shape
    assert rel_err(aux["semantics"].cpu(), g["semantics_map"]) < TOL


def test_sampler_normals_vs_oracle(nfi):
    """sampler(..., ['sigma', 'rgb', 'normals']) (generator.py:599-623) at explicit points, some of them
    outside the box / on its faces, against the CPU oracle."""
    import nfi_oracle as O
    g = load("decoder_sdf_attn")
    B, res = g["planes"].shape[0], g["planes"].shape[-1]
    fp = O.FieldParams(g["W0"], g["b0"], g["W1"], g["b1"], 1.4, True, 10, g["beta"], g["alpha"])
    sig0, rgb0, _, nrm0 = O.field_eval(g["planes"].view(B, 3, 32, res, res), g["palette"], g["pts"], fp,
                                       want_normals=True)
    fld = _field(nfi, g, 10, True)
    smp = nfi.dec.make_sampler(cu(g["planes"]), cu(g["palette"]), fld)
    out = smp(cu(g["pts"]).unsqueeze(2), ["sigma", "rgb", "normals"])
    assert not out["sigma"].requires_grad
    assert rel_err(out["sigma"].cpu(), sig0) < TOL and rel_err(out["rgb"].cpu(), rgb0) < TOL
    # compare un-normalised directions where the gradient is not tiny; everywhere: unit or zero length
    n1 = out["normals"].cpu().reshape(nrm0.shape)
    ln = n1.norm(dim=-1)
    assert bool((((ln - 1).abs() < 1e-5) | (ln == 0)).all())
    assert rel_err(n1, nrm0) < NORMALS_TOL


# ---------------------------------------------------------------- fused C-ABI render, CUDA graph
def _crender_case(nfi, B=2, HW=32, perspective=True, white=False, seed=21):
    from nfi_b200 import synth
    H = W = HW
    S, res = 64, 64
    planes = synth.make_planes(B, res=res, seed=1 + seed)
    fd = synth.make_field(planes, seed=3 + seed)
    pal = synth.make_palette(B, 10, seed=4 + seed)
    cam, focal = synth.make_cameras(B, seed=2 + seed, perspective=perspective)
    g = torch.Generator().manual_seed(seed)
    u_s, u_p = torch.rand(B, H, W, S, generator=g).cuda(), torch.rand(B * H * W, S, generator=g).cuda()
    target = (torch.rand(B, H, W, 3, generator=g) * 2 - 1).cuda()
    gmask = (torch.rand(B, H, W, generator=g) - 0.5).cuda()
    fld = nfi.dec.Field(*[fd[k].cuda() for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], 1.4, 10, True)
    return dict(B=B, H=H, W=W, S=S, planes=planes.cuda(), pal=pal.cuda(), cam=cam.cuda(),
                focal=focal.cuda() if perspective else None, u_s=u_s, u_p=u_p, target=target, gmask=gmask,
                fld=fld, white=white)


@pytest.mark.parametrize("perspective,white", [(True, False), (False, True)])
def test_fused_c_render_matches_autograd_path(nfi, perspective, white):
    """nfi_render_fwd / nfi_render_bwd (one C call each, one workspace) against render_planes + autograd:
    the same kernels in the same order, so images are bit-identical; gradients differ only by the order
    of the fp32 atomics."""
    from nfi_b200 import crender
    c = _crender_case(nfi, perspective=perspective, white=white)
    lv = [t.clone().requires_grad_() for t in (c["planes"], c["pal"], c["cam"])] + \
         ([c["focal"].clone().requires_grad_()] if perspective else [])
    focal = lv[3] if perspective else None
    rgb, depth, mask = nfi.render.render_planes(lv[0], lv[1], c["fld"], c["H"], c["W"], lv[2], focal, None, None,
                                                c["S"], randomize=True, white_background=white,
                                                u_strat=c["u_s"], u_pdf=c["u_p"])
    loss = ((rgb - c["target"]).abs().sum() + (mask * c["gmask"]).sum())
    g_rgb, g_mask = torch.autograd.grad(loss, [rgb, mask], retain_graph=True)
    grads = torch.autograd.grad(loss, lv)
    pcl = nfi.ops.planes_to_channels_last(c["planes"])
    cr = crender.CRender(pcl, c["pal"], c["fld"], c["H"], c["W"], c["cam"], c["focal"], None, None, c["S"],
                         randomize=True, white_background=white, u_strat=c["u_s"], u_pdf=c["u_p"])
    rgb2, depth2, mask2 = cr.forward()
    assert torch.equal(rgb2, rgb.detach()) and torch.equal(depth2, depth) and torch.equal(mask2, mask.detach())
    out = cr.backward(g_rgb, g_mask)
    g_planes_cl = nfi.ops.planes_to_channels_last(grads[0])         # same layout as the C path's output
    assert rel_err(out["planes_cl"].cpu(), g_planes_cl.cpu()) < 1e-5
    assert rel_err(out["palette"].cpu(), grads[1].cpu()) < 1e-5
    assert rel_err(out["cam2world"].cpu(), grads[2].cpu()) < 1e-5
    if perspective:
        assert rel_err(out["focal"].cpu(), grads[3].cpu()) < 1e-5


def test_fused_c_render_philox_and_errors(nfi):
    """No uniforms supplied: both passes draw Philox numbers from (seed, offset) -- reproducible per seed,
    different across seeds, image statistics close to the torch.rand render; malformed calls come back as
    NfiError with the library's message."""
    from nfi_b200 import crender, _lib
    c = _crender_case(nfi, B=2, HW=32)
    pcl = nfi.ops.planes_to_channels_last(c["planes"])
    mk = lambda seed: crender.CRender(pcl, c["pal"], c["fld"], c["H"], c["W"], c["cam"], c["focal"], None, None,
                                      c["S"], randomize=True, seed=seed, with_backward=False).forward()
    a, b, d = mk(7), mk(7), mk(8)
    assert torch.equal(a[0], b[0]) and not torch.equal(a[0], d[0])
    ref = nfi.render.render_planes(c["planes"], c["pal"], c["fld"], c["H"], c["W"], c["cam"], c["focal"], None, None,
                                   c["S"], randomize=True, u_strat=c["u_s"], u_pdf=c["u_p"])
    assert abs(float(a[2].mean()) - float(ref[2].mean())) < 0.02
    assert float((a[0] - ref[0]).abs().mean()) < 0.05
    with pytest.raises(_lib.NfiError):
        crender.CRender(pcl, None, c["fld"], c["H"], c["W"], c["cam"], c["focal"], None, None, c["S"]).forward()


def test_graphed_step_matches_eager(nfi):
    """nfi_b200.graph.GraphedStep: the captured render forward + loss + backward replays to the eager
    results, and new input values are picked up from the static buffers."""
    from nfi_b200 import graph
    c = _crender_case(nfi, B=2, HW=32)

    def step(planes, pal, cam, focal, target):
        rgb, depth, mask = nfi.render.render_planes(planes, pal, c["fld"], c["H"], c["W"], cam, focal, None, None,
                                                    c["S"], randomize=True, u_strat=c["u_s"], u_pdf=c["u_p"],
                                                    check_hits="defer")
        loss = (rgb - target).abs().mean()
        return (loss, rgb, *torch.autograd.grad(loss, [planes, pal, cam, focal]))

    leaves = [t.clone().requires_grad_() for t in (c["planes"], c["pal"], c["cam"], c["focal"])] + [c["target"]]
    gs = graph.GraphedStep(step, leaves)
    for trial in range(2):
        if trial == 1:      # perturbed inputs go through the same graph
            leaves = [(t.detach() * 1.01).requires_grad_(t.requires_grad) for t in leaves]
        ref = step(*leaves)
        out = gs(*leaves)
        gs.check_hits()
        assert torch.equal(out[0], ref[0].detach()) and torch.equal(out[1], ref[1].detach())
        for a, b in zip(out[2:], ref[2:]):
            assert rel_err(a.cpu(), b.cpu()) < 1e-5
    assert gs.launches_per_replay == 16


def test_graphed_host_pipeline_matches_eager_pipeline(nfi):
    """host_pipeline.GraphedHostStepPipeline (chunk step replayed as CUDA graphs, two instances used
    round-robin) against the eager HostStepPipeline on the same pinned host buffers; two runs, so the
    buffers of both graph instances are reused."""
    from nfi_b200 import host_pipeline
    B, n_chunks = 4, [1, 2, 1]                 # unequal chunks: two graph sizes
    c = _crender_case(nfi, B=B, HW=32)
    H, W, S = c["H"], c["W"], c["S"]

    def step(planes, pal, cam, focal, target, u_s, u_p):
        planes, pal, cam, focal = [t.requires_grad_() for t in (planes, pal, cam, focal)]
        rgb, depth, mask = nfi.render.render_planes(planes, pal, c["fld"], H, W, cam, focal, None, None, S,
                                                    randomize=True, u_strat=u_s, u_pdf=u_p, check_hits="defer")
        loss = (rgb - target).abs().sum()
        return (loss, rgb, depth, mask, *torch.autograd.grad(loss, [planes, pal, cam, focal]))

    pin = lambda t: t.detach().cpu().contiguous().pin_memory()
    host_in = [pin(t) for t in (c["planes"], c["pal"], c["cam"], c["focal"], c["target"], c["u_s"],
                                c["u_p"].view(B, H * W, S))]
    step_v = lambda p, a, cm, f, tg, us, up: step(p, a, cm, f, tg, us, up.reshape(-1, S))
    shapes = ((), (B, H, W, 3), (B, H, W), (B, H, W), tuple(c["planes"].shape), tuple(c["pal"].shape), (B, 4, 4), (B,))
    out_e = [torch.zeros(sh).pin_memory() for sh in shapes]
    out_g = [torch.zeros(sh).pin_memory() for sh in shapes]
    eager = host_pipeline.HostStepPipeline(step_v, B, n_chunks, torch.device("cuda", 0))
    graphed = host_pipeline.GraphedHostStepPipeline(step_v, B, n_chunks, torch.device("cuda", 0), host_in, depth=2)
    assert sorted(graphed.launches_per_chunk) == [1, 2] and min(graphed.launches_per_chunk.values()) > 0
    for trial in range(2):
        if trial == 1:
            host_in[0].mul_(1.01)
            host_in[4].mul_(0.9)
        eager.run(host_in, out_e)
        graphed.run(host_in, out_g)
        for k in range(4):
            assert torch.equal(out_e[k], out_g[k]) if k else rel_err(out_g[0], out_e[0]) < 1e-6, k
        for k in range(4, 8):
            assert rel_err(out_g[k], out_e[k]) < 1e-5, k
        assert float(out_g[2].abs().max()) > 0


# ---------------------------------------------------------------- the paths the fast kernels fall back from
@pytest.mark.parametrize("B,H,W,S,res", [
    (1, 10, 14, 32, 20),     # ragged image (linear tile order), row-major texels (res % 8 != 0), general merge
    (2, 8, 16, 30, 24),      # S % 4 != 0: scalar stratified kernel, partial rows in the scans
    (1, 16, 16, 64, 36),     # S = 64 fast merge with the scalar plane transpose (HW % 128 != 0)
])
def test_render_vs_oracle_fallback_paths(nfi, B, H, W, S, res):
    _check(*_oracle_vs_cuda(nfi, B, H, W, S, res, True, False, False, False, seed=3))


def test_merge_fast_path_ties_and_unsorted_coarse(nfi):
    """merge_composite's S = 64 fast path must hand rays with exactly equal fine depths, fine == coarse
    ties and a non-ascending coarse list to the general path: permutation bit-exact against
    torch.sort(stable=True) of cat(coarse, fine) (run.py:283-285) in every case."""
    g = torch.Generator().manual_seed(5)
    R, S = 64, 64
    zc = torch.sort(torch.rand(R, S, generator=g) * 2 + 0.5, dim=-1).values
    zf = torch.rand(R, S, generator=g) * 2 + 0.5
    zf[1, 5] = zf[1, 40]                       # two equal fine depths
    zf[2, 7] = zc[2, 20]                       # fine == coarse: coarse first
    zc[3, 10], zc[3, 11] = zc[3, 11].clone(), zc[3, 10].clone()     # coarse not ascending
    zf[4] = zf[4, 0]                           # all fine depths equal
    zc[5] = zc[5, 0]; zf[5] = zc[5, 0]         # everything equal
    coarse = torch.rand(R, S, 4, generator=g); fine = torch.rand(R, S, 4, generator=g)
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    rgb, depth, mask, perm, zs = nfi.ops.merge_composite(cu(coarse), cu(fine), cu(zc), cu(zf), cu(d), False)
    ref = torch.sort(torch.cat((zc, zf), dim=-1), dim=-1, stable=True)
    assert torch.equal(perm.cpu().long(), ref.indices)
    assert torch.equal(zs.cpu(), ref.values)


# ---------------------------------------------------------------- SURVEY.md 8(f) N3
@pytest.mark.parametrize("name,persp,flipped", [("persp_flip", True, True), ("persp", True, False),
                                                ("ortho_flip", False, True)])
def test_pose_to_matrix_golden(nfi, name, persp, flipped):
    """F.normalize(q) + pose_utils.pose_to_matrix (run.py:2257-2262) against the unmodified reference."""
    from nfi_b200 import pose_utils
    g = {k[len(name) + 1:]: v for k, v in load("pose").items() if k.startswith(name + "_")}
    if name == "persp":
        g = {k: v for k, v in g.items() if not k.startswith("flip_")}
    z0 = cu(g["z0"]).requires_grad_() if persp else None
    t2, s, q = [cu(g[k]).requires_grad_() for k in ("t2", "s", "q")]
    mat, focal = pose_utils.pose_to_matrix(z0, t2, s, q, flipped, normalize_quaternion=True)
    assert rel_err(mat.cpu(), g["mat"]) < 1e-6
    loss = (mat * cu(g["gm"])).sum()
    if persp:
        assert rel_err(focal.cpu(), g["focal"]) < 1e-6
        loss = loss + (focal * cu(g["gf"])).sum()
    else:
        assert focal is None
    leaves = ([z0] if persp else []) + [t2, s, q]
    keys = (["d_z0"] if persp else []) + ["d_t2", "d_s", "d_q"]
    for a, k in zip(torch.autograd.grad(loss, leaves), keys):
        assert rel_err(a.cpu(), g[k]) < 1e-5, k
    # the plain signature (already-normalised quaternion) agrees with the fused one
    qn = torch.nn.functional.normalize(q.detach(), dim=-1)
    mat2, _ = pose_utils.pose_to_matrix(z0, t2, s, qn, flipped)
    assert rel_err(mat2.cpu(), g["mat"]) < 1e-6
    if not persp:
        with pytest.raises(RuntimeError):
            pose_utils.pose_to_matrix(None, t2.repeat(2, 1), s.repeat(2), q.repeat(2, 1), flipped)
