"""Property tests of the fused render, driven by `hypothesis` over image sizes,
camera models, seeds, background and sample counts.  Every call goes through the C ABI."""
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu
SET = dict(max_examples=12, deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)


def _scene(B, res, seed, perspective):
    from nfi_b200 import decoder, synth
    planes = synth.make_planes(B, res=res, seed=1 + seed)
    fd = synth.make_field(planes, seed=3 + seed)
    fld = decoder.Field(*[fd[k].cuda() for k in ("W0", "b0", "W1", "b1", "beta", "alpha")], 1.4, 10, True)
    pal = synth.make_palette(B, 10, seed=4 + seed).cuda()
    cam, focal = synth.make_cameras(B, seed=2 + seed, perspective=perspective)
    return planes.cuda(), fld, pal, cam.cuda(), (focal.cuda() if perspective else None)


def _render(planes, fld, pal, cam, focal, H, W, S, white, u_s, u_p, **kw):
    from nfi_b200 import render
    return render.render_planes(planes, pal, fld, H, W, cam, focal, None, None, S, randomize=True,
                                white_background=white, u_strat=u_s, u_pdf=u_p, return_aux=True, **kw)


@settings(**SET)
@given(hb=st.integers(1, 4), wb=st.integers(1, 3), B=st.integers(1, 3), seed=st.integers(0, 50),
       perspective=st.booleans(), white=st.booleans(), S=st.sampled_from([16, 32, 64]))
def test_images_are_bounded_and_samples_sorted(hb, wb, B, seed, perspective, white, S):
    H, W = 8 * hb, 16 * wb                                   # the pixel-block tile schedule
    planes, fld, pal, cam, focal = _scene(B, 32, seed, perspective)
    g = torch.Generator(device="cuda").manual_seed(seed)
    u_s = torch.rand(B, H, W, S, device="cuda", generator=g)
    u_p = torch.rand(B * H * W, S, device="cuda", generator=g)
    rgb, depth, mask, aux = _render(planes, fld, pal, cam, focal, H, W, S, white, u_s, u_p)
    assert torch.isfinite(rgb).all() and torch.isfinite(depth).all() and torch.isfinite(mask).all()
    # mask = sum of the compositing weights (lib/nerf_utils.py:156): in [0, 1]
    assert float(mask.min()) >= 0.0 and float(mask.max()) <= 1.0 + 1e-5
    # merged depths ascend (run.py:283-285) and are a permutation of cat(coarse, fine)
    zs = aux["z_sorted"].reshape(-1, 2 * S)
    assert bool((zs[:, 1:] >= zs[:, :-1]).all())
    both = torch.cat((aux["z_coarse"].reshape(-1, S), aux["z_fine"].reshape(-1, S)), -1)
    assert torch.equal(torch.sort(both, -1)[0], zs)
    assert torch.equal(torch.sort(aux["perm"].reshape(-1, 2 * S).long(), -1)[0],
                       torch.arange(2 * S, device="cuda").expand(zs.shape[0], -1))
    # rgb = sum_i w_i c_i (+ (1 - mask) for a white background), c_i a convex combination of the image's
    # palette rows (softmax . attention values, models/generator.py:668-678): inside the hull scaled by mask
    lo = pal.amin(1).view(B, 1, 1, 3)
    hi = pal.amax(1).view(B, 1, 1, 3)
    base = (1.0 - mask).unsqueeze(-1) if white else 0.0
    assert bool((rgb <= hi * mask.unsqueeze(-1) + base + 1e-4).all())
    assert bool((rgb >= lo * mask.unsqueeze(-1) + base - 1e-4).all())


@settings(**SET)
@given(B=st.integers(2, 4), seed=st.integers(0, 50), perspective=st.booleans())
def test_batch_permutation_invariance(B, seed, perspective):
    """Images are independent: permuting the batch permutes the outputs, bit for bit."""
    H, W, S = 16, 32, 32
    planes, fld, pal, cam, focal = _scene(B, 32, seed, perspective)
    g = torch.Generator(device="cuda").manual_seed(seed)
    u_s = torch.rand(B, H, W, S, device="cuda", generator=g)
    u_p = torch.rand(B, H * W, S, device="cuda", generator=g)
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(seed)).cuda()
    a = _render(planes, fld, pal, cam, focal, H, W, S, False, u_s, u_p.reshape(-1, S))
    b = _render(planes[perm].contiguous(), fld, pal[perm].contiguous(), cam[perm].contiguous(),
                focal[perm].contiguous() if perspective else None, H, W, S, False, u_s[perm].contiguous(),
                u_p[perm].reshape(-1, S).contiguous())
    # (the miss-ray near/far fill is a min / max over the whole batch, lib/nerf_utils.py:260-261: the same set
    # of rays either way)
    for x, y in zip(a[:3], b[:3]):
        assert torch.equal(x[perm], y)


@settings(**SET)
@given(seed=st.integers(0, 50), white=st.booleans())
def test_rays_that_miss_the_box_return_the_background(seed, white):
    """A camera far off-axis: part of the image misses the scene box; those pixels carry no density
    (the out-of-box mask, models/generator.py:605-607) -> mask 0, rgb = background."""
    from nfi_b200 import ops
    H, W, S, B = 16, 32, 32, 1
    planes, fld, pal, cam, focal = _scene(B, 32, seed, True)
    focal = focal * 0.45                                      # wide field of view: the box does not fill the frame
    o, d = ops.ray_bundle(H, W, focal, cam, None, None, normalize=True)
    inv = 1.0 / d
    t0, t1 = (-1.4 - o) * inv, (1.4 - o) * inv
    lo_t, hi_t = torch.minimum(t0, t1).amax(-1), torch.maximum(t0, t1).amin(-1)
    hit = lo_t <= hi_t
    g = torch.Generator(device="cuda").manual_seed(seed)
    u_s = torch.rand(B, H, W, S, device="cuda", generator=g)
    u_p = torch.rand(B * H * W, S, device="cuda", generator=g)
    rgb, depth, mask, _ = _render(planes, fld, pal, cam, focal, H, W, S, white, u_s, u_p)
    miss = lo_t > hi_t + 1e-3                                # clear misses only (grazing rays are a rounding matter)
    if int(miss.sum()) == 0 or int(hit.sum()) == 0:
        return
    assert float(mask[miss].abs().max()) == 0.0
    assert float((rgb[miss] - (1.0 if white else 0.0)).abs().max()) == 0.0


def test_composite_gradient_matches_fp64_finite_differences():
    """Gradient check of A10 (lib/nerf_utils.py:125-163) against central differences of an fp64 restatement."""
    from nfi_b200 import ops
    R, S = 6, 16
    g = torch.Generator().manual_seed(3)
    sigma = torch.rand(R, S, generator=g) * 3
    rgbs = torch.rand(R, S, 3, generator=g) * 2 - 1
    z = torch.sort(torch.rand(R, S, generator=g) * 2 + 1, -1)[0]
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    w_rgb, w_mask = torch.randn(R, 3, generator=g), torch.randn(R, generator=g)

    def loss64(sig, col):
        dist = torch.cat((z[:, 1:] - z[:, :-1], torch.zeros(R, 1)), -1).double() * d.double().norm(dim=-1, keepdim=True)
        alpha = 1 - torch.exp(-sig * dist)
        T = torch.cumprod(torch.cat((torch.ones(R, 1, dtype=torch.float64), 1 - alpha + 1e-10), -1), -1)[:, :-1]
        w = alpha * T
        return ((w[..., None] * col).sum(1) * w_rgb.double()).sum() + (w.sum(1) * w_mask.double()).sum()

    s_c, c_c = sigma.cuda().requires_grad_(), rgbs.cuda().requires_grad_()
    rgb, depth, mask, _, _ = ops.composite(s_c, c_c, d.cuda(), z.cuda(), None, None, False)
    ((rgb * w_rgb.cuda()).sum() + (mask * w_mask.cuda()).sum()).backward()
    s64, c64 = sigma.double(), rgbs.double()
    eps = 1e-6
    for (r, k) in [(0, 0), (1, 5), (2, 15), (3, 7), (5, 14)]:
        e = torch.zeros_like(s64); e[r, k] = eps
        fd = float(loss64(s64 + e, c64) - loss64(s64 - e, c64)) / (2 * eps)
        assert abs(float(s_c.grad[r, k]) - fd) <= 1e-4 * max(1.0, abs(fd)), (r, k, float(s_c.grad[r, k]), fd)
        e3 = torch.zeros_like(c64); e3[r, k, 1] = eps
        fd = float(loss64(s64, c64 + e3) - loss64(s64, c64 - e3)) / (2 * eps)
        assert abs(float(c_c.grad[r, k, 1]) - fd) <= 1e-4 * max(1.0, abs(fd))
