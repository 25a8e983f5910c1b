"""Pins oracle/nfi_oracle.py to the LIVE reference (only where /root/reference is mounted).

The same comparisons against *stored* reference outputs run everywhere in
tests/test_oracle_golden.py.
"""
import types

import pytest
import torch

import nfi_oracle as O
import ref_loader
from nfi_b200 import synth

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference checkout absent")


def eq(a, b):
    """Bitwise-equal values (NaN == NaN: the slab test produces NaN for rays grazing a face)."""
    return a.shape == b.shape and bool(((a == b) | (a.isnan() & b.isnan())).all())


def cams(kind, B=2):
    if kind == "persp":
        cam, focal = synth.make_cameras(B, seed=2)
        return cam, focal
    if kind == "ortho":
        cam, _ = synth.make_cameras(B, seed=5, perspective=False)
        return cam, None
    return synth.make_ortho_cub_cameras(B, seed=7), None


@pytest.mark.parametrize("kind", ["persp", "ortho", "cub"])
@pytest.mark.parametrize("with_bbox", [False, True])
@pytest.mark.parametrize("with_center", [False, True])
def test_pixel_rays_bitexact(kind, with_bbox, with_center):
    nu = ref_loader.modules().nerf_utils
    cam, focal = cams(kind, 3)
    g = torch.Generator().manual_seed(11)
    bbox = None
    if with_bbox:
        bbox = torch.stack((torch.rand(3, 2, generator=g) - 0.5, 0.5 + torch.rand(3, 2, generator=g)), 1)
    center = 0.5 + (torch.rand(3, 2, generator=g) - 0.5) * 0.6 if with_center else None
    o_ref, d_ref = nu.get_ray_bundle(24, 40, focal, cam, bbox, center)
    o, d = O.pixel_rays(24, 40, cam, focal, center, bbox)
    assert eq(o, o_ref.expand_as(o)) and eq(d, d_ref)
    assert eq(O.unit_directions(d), torch.nn.functional.normalize(d_ref, dim=-1))


def test_pixel_rays_scalar_focal():
    nu = ref_loader.modules().nerf_utils
    cam, focal = cams("persp", 1)
    f0 = focal[0].clone()                       # 0-dim, as in eval_kitti_persp.py:948
    center = torch.tensor([[0.61, 0.43]])
    o_ref, d_ref = nu.get_ray_bundle(16, 16, f0, cam, None, center)
    o, d = O.pixel_rays(16, 16, cam, f0, center, None)
    assert eq(d, d_ref)


@pytest.mark.parametrize("kind", ["persp", "ortho", "cub"])
def test_near_far_bitexact(kind):
    nu = ref_loader.modules().nerf_utils
    cam, focal = cams(kind, 2)
    o, d = O.pixel_rays(32, 32, cam, focal)
    d = O.unit_directions(d)
    r = 2.0 if kind == "cub" else 1.0      # small box -> some rays miss
    n_ref, f_ref = nu.compute_near_far_planes(o, d, r)
    n, f, hit = O.box_near_far(o, d, r, return_hit=True)
    assert eq(n, n_ref) and eq(f, f_ref)
    if kind == "persp":
        assert 0 < hit.float().mean() < 1


def test_near_far_axis_aligned_and_all_miss():
    nu = ref_loader.modules().nerf_utils
    o = torch.tensor([[0.0, 0.0, 3.0], [0.2, -0.3, 3.0], [5.0, 0.0, 3.0], [1.4, 0.0, 3.0]])
    d = torch.tensor([[0.0, 0.0, -1.0], [0.0, -0.0, -1.0], [0.0, 0.0, -1.0], [0.0, 0.0, -1.0]])
    n_ref, f_ref = nu.compute_near_far_planes(o, d, 1.4)
    n, f = O.box_near_far(o, d, 1.4)
    assert eq(n, n_ref) and eq(f, f_ref)
    with pytest.raises(RuntimeError):
        O.box_near_far(o[2:3], d[2:3], 1.4)
    with pytest.raises(RuntimeError):
        nu.compute_near_far_planes(o[2:3], d[2:3], 1.4)


@pytest.mark.parametrize("randomize", [False, True])
def test_stratified_bitexact(randomize):
    nu = ref_loader.modules().nerf_utils
    cam, focal = cams("persp", 2)
    o, d = O.pixel_rays(16, 16, cam, focal)
    d = O.unit_directions(d)
    near, far = O.box_near_far(o, d, 1.4)
    torch.manual_seed(5)
    q_ref, z_ref = nu.compute_query_points_from_rays(o, d, near, far, 64, randomize=randomize)
    torch.manual_seed(5)
    u = torch.rand(2, 16, 16, 64) if randomize else None
    z = O.stratified_depths(near, far, 64, u)
    assert eq(z, z_ref)
    assert eq(O.points_on_rays(o, d, z), q_ref)


def test_weights_smoothing_pdf_bitexact():
    nu = ref_loader.modules().nerf_utils
    F = torch.nn.functional
    g = torch.Generator().manual_seed(3)
    R, S = 4096, 64
    o = torch.zeros(R, 3)
    d = F.normalize(torch.randn(R, 3, generator=g), dim=-1)
    near = 0.5 + torch.rand(R, generator=g)
    far = near + 0.5 + 2 * torch.rand(R, generator=g)
    z = O.stratified_depths(near, far, S, torch.rand(R, S, generator=g))
    sigma = torch.rand(R, S, generator=g) * (torch.rand(R, S, generator=g) > 0.6) * 30
    w_ref = nu.render_volume_density_weights_only(sigma, o, d, z)
    w = O.transmittance_weights(sigma, z, O.vec3_norm(d))
    assert eq(w, w_ref)
    ws_ref = F.avg_pool1d(F.max_pool1d(w_ref.unsqueeze(1), 2, 1, padding=1), 2, 1).squeeze() + 0.01
    assert eq(O.smooth_weights(w), ws_ref)
    mid = 0.5 * (z[..., 1:] + z[..., :-1])
    for det in (True, False):
        torch.manual_seed(9)
        zs_ref = nu.sample_pdf(mid, ws_ref[..., 1:-1], S, deterministic=det)
        torch.manual_seed(9)
        u = None if det else torch.rand(R, S)
        zs = O.inverse_cdf_samples(mid, ws_ref[..., 1:-1], S, u)
        assert eq(zs, zs_ref), (det, (zs != zs_ref).float().mean())


def test_composite_matches():
    nu = ref_loader.modules().nerf_utils
    g = torch.Generator().manual_seed(4)
    R, S = 512, 128
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    z = torch.sort(torch.rand(R, S, generator=g) * 3 + 0.2, dim=-1).values
    sigma = torch.rand(R, S, generator=g) * 20
    rgb = torch.rand(R, S, 3, generator=g) * 2 - 1
    sem = torch.rand(R, S, 10, generator=g)
    for wb in (False, True):
        ref = nu.render_volume_density(sigma, rgb, torch.zeros(R, 3), d, z, None, sem, white_background=wb)
        out = O.composite(sigma, rgb, z, O.vec3_norm(d), wb, semantics=sem)
        for a, b in zip(out, ref):
            if b is not None:
                torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-6)


class _PlanesStub(torch.nn.Module):
    """Stands in for the StyleGAN2 synthesis network so the reference Generator runs on given planes."""

    def __init__(self, planes):
        super().__init__()
        self.planes = planes

    def forward(self, ws, **kw):
        return self.planes


def build_reference_generator(planes, field, scene_range, A=10, use_sdf=True):
    gen_mod = ref_loader.modules().generator
    torch.manual_seed(0)
    G = gen_mod.Generator(512, scene_range, attention_values=A, use_viewdir=False, use_encoder=False,
                          disable_stylegan_noise=True, use_sdf=use_sdf).eval()
    G.synthesis_network = _PlanesStub(planes)
    with torch.no_grad():
        G.decoder.net[0].weight.copy_(field["W0"]); G.decoder.net[0].bias.copy_(field["b0"])
        G.decoder.net[2].weight.copy_(field["W1"]); G.decoder.net[2].bias.copy_(field["b1"])
        if use_sdf:
            G.beta.copy_(field["beta"]); G.alpha.copy_(field["alpha"])
    return G.requires_grad_(False)


@pytest.mark.parametrize("case", ["persp_rand", "ortho_det", "persp_white_nocam"])
def test_render_matches_lifted_reference(case):
    B, H, W, S, res = 2, 12, 12, 64, 32
    persp = case.startswith("persp")
    randomize = case != "ortho_det"
    white = "white" in case
    scene_range = 1.4
    planes = synth.make_planes(B, res, seed=1).requires_grad_()
    field = synth.make_field(planes.detach(), seed=3)
    palette = synth.make_palette(B, 10, seed=4).requires_grad_()
    z0, t2, s, q = synth.make_pose_params(B, seed=2, perspective=persp)
    leaves = [t.requires_grad_() for t in (t2, s, q)] + ([z0.requires_grad_()] if persp else [])
    cam, focal = synth.pose_to_matrix(z0, t2, s, q, True)
    # also pins synth.pose_to_matrix to lib/pose_utils.py:48-75
    # (perspective only: the reference's orthographic branch divides [B,3] by [B] and only
    #  broadcasts for B in {1, 3} -- lib/pose_utils.py:69)
    if persp:
        cam_ref, focal_ref = ref_loader.modules().pose_utils.pose_to_matrix(z0, t2, s, q, True)
        torch.testing.assert_close(cam, cam_ref, rtol=1e-6, atol=1e-6)
    G = build_reference_generator(planes, field, scene_range)
    render_ref, _ = ref_loader.lift_render(scene_range=scene_range, white_background=white)
    ws = torch.zeros(B, 15, 512)
    no_cam = "nocam" in case
    torch.manual_seed(77)
    out_ref = render_ref(G, H, W, cam, focal, None, None, ws, S, randomize=randomize,
                         extra_model_inputs={"attention_values": palette}, force_no_cam_grad=no_cam)
    torch.manual_seed(77)
    u1 = torch.rand(B, H, W, S) if randomize else None
    u2 = torch.rand(B * H * W, S) if randomize else None
    fp = O.FieldParams(field["W0"], field["b0"], field["W1"], field["b1"], scene_range, True, 10,
                       field["beta"], field["alpha"])
    out = O.render(planes.view(B, 3, 32, res, res), palette, fp, H, W, cam, focal, S=S,
                   white_background=white, u_strat=u1, u_pdf=u2, randomize=randomize,
                   no_cam_grad=no_cam)
    for a, b in zip(out, out_ref[:3]):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
    assert out_ref[2].mean() > 0.2          # the fixture is not an empty image
    tgt = torch.rand(B, H, W, 3, generator=torch.Generator().manual_seed(8)) * 2 - 1
    wants = [planes, palette] + ([] if no_cam else leaves)
    g_ref = torch.autograd.grad((out_ref[0] - tgt).abs().mean() + out_ref[2].mean(), wants,
                                retain_graph=True, allow_unused=True)
    g = torch.autograd.grad((out[0] - tgt).abs().mean() + out[2].mean(), wants, allow_unused=True)
    for a, b in zip(g, g_ref):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-7)
