"""CPU oracle for the nerf-from-image volume-rendering hot path -- TEST INFRASTRUCTURE ONLY.

This file is a *restatement* (written from scratch, flat-ray formulation, fp32 on
the CPU) of the algorithm in the reference's hot path.  It is the checker the
CUDA kernels are compared against.  It is NOT a product code path:

  * only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
    `--impl reference` leg may import it;
  * nothing under `nerf-from-image_b200/` imports it, and the product fails
    loudly when its CUDA library is missing instead of falling back here.

Parity status: PINNED.  `oracle/make_golden.py` runs the unmodified reference
(`/root/reference`, imported, `run.py::render` lifted with `ast`) in the build
container and stores its inputs/outputs under `tests/golden/*.npz`;
`tests/test_oracle_golden.py` checks every function below against those
vectors (and, when the reference checkout is present, against the live
reference in `tests/test_oracle_vs_reference.py`).

Third-party arithmetic: everything is PyTorch ATen (torch 2.11.0).  Semantics
that live in ATen rather than in the reference source were probed and are
restated explicitly where they matter for the bit-exact index contract:

  * `torch.lerp(a, b, w)`         = w < 0.5 ? fma(w, b-a, a) : fma(w-1, b-a, b)
  * `x.norm(p=2, dim=-1)` (3-vec) = sqrt(fma(c, c, fma(b, b, a*a)))
  * `tensor / python_float`       = IEEE division (not multiply-by-reciprocal)
  * `cumsum` / `cumprod` (CPU)    = double accumulator, each output rounded to fp32
  * `sum(dim=-1)` (CPU, contiguous row) = 8-lane vector partials, 4-way ILP, see
    `aten_row_sum` -- restated with explicit adds so the oracle gives the same
    bits on any host CPU.

Reference citations are `file:line` relative to the reference checkout.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn.functional as F

f32 = torch.float32


# --------------------------------------------------------------------------------------
# ATen reduction order restated (see module docstring)
# --------------------------------------------------------------------------------------
def aten_row_sum(x: torch.Tensor) -> torch.Tensor:
    """fp32 sum over the last (contiguous) dim in ATen's CPU order (SumKernel, Vec=8, ILP=4).

    Valid for rows shorter than 512 elements (no cascade level is reached).  Used for
    `weights.sum(-1)` at lib/nerf_utils.py:192 and :156 so that the CDF / mask are
    reproducible bit for bit on any machine.
    """
    n = x.shape[-1]
    V, ILP = 8, 4
    nv = n // V
    vec = [x[..., i * V:(i + 1) * V] for i in range(nv)]
    rows = nv // ILP
    part = [None] * ILP
    for r in range(rows):
        for k in range(ILP):
            v = vec[r * ILP + k]
            part[k] = v if part[k] is None else part[k] + v
    zero = torch.zeros(x.shape[:-1] + (V,), dtype=x.dtype, device=x.device)
    part = [zero if p is None else p for p in part]
    for i in range(rows * ILP, nv):
        part[0] = part[0] + vec[i]
    for k in range(1, ILP):
        part[0] = part[0] + part[k]
    acc = torch.zeros(x.shape[:-1], dtype=x.dtype, device=x.device)
    for k in range(nv * V, n):
        acc = acc + x[..., k]
    for k in range(V):
        acc = acc + part[0][..., k]
    return acc


def cumsum_f64(x):
    return torch.cumsum(x.double(), dim=-1).to(x.dtype)


def cumprod_f64(x):
    return torch.cumprod(x.double(), dim=-1).to(x.dtype)


def vec3_norm(v):
    """`v.norm(p=2, dim=-1)` for 3-vectors in ATen's fused order."""
    a, b, c = v[..., 0].double(), v[..., 1].double(), v[..., 2].double()
    acc = (a * a).float()
    acc = (b * b + acc.double()).float()   # fma: exact product, single rounding
    acc = (c * c + acc.double()).float()
    return acc.double().sqrt().float()   # torch.sqrt(fp32) on CPU is not always correctly rounded


# --------------------------------------------------------------------------------------
# A1  pixel rays                      lib/nerf_utils.py:28-93  (+ F.normalize, run.py:196)
# --------------------------------------------------------------------------------------
def pixel_rays(H: int, W: int, cam2world, focal=None, center=None, bbox=None):
    """Ray origins/directions [B,H,W,3]; perspective iff `focal` is given.

    Pixel (h, w) of image b is ray index (b*H + h)*W + w -- the bit-exact ray order
    contract.  `bbox[b] = [[bx, by], [bw, bh]]`, `center[b] = [cx, cy]`.
    """
    B = cam2world.shape[0]
    dev = cam2world.device                    # CPU is the pinned oracle; CUDA = the reference's eager GPU path
    u = (torch.arange(W, dtype=torch.int64, device=dev) / W).to(f32).view(1, 1, W).expand(1, H, W)
    v = (torch.arange(H, dtype=torch.int64, device=dev) / H).to(f32).view(1, H, 1).expand(1, H, W)
    R = cam2world[:, :3, :3]                  # [B,3,3]
    t = cam2world[:, :3, 3]                   # [B,3]
    if focal is not None:                     # nerf_utils.py:40-67
        if center is not None:
            u = u - 0.5 * (2 * center[:, 0].view(B, 1, 1) - 1) - 0.5
            v = v - 0.5 * (2 * center[:, 1].view(B, 1, 1) - 1) - 0.5
        else:
            u = u - 0.5
            v = v - 0.5
        if bbox is not None:
            u = (bbox[:, 1, 0].view(B, 1, 1) * (u + 0.5) + bbox[:, 0, 0].view(B, 1, 1)) * 0.5
            v = -(bbox[:, 1, 1].view(B, 1, 1) * (-v + 0.5) + bbox[:, 0, 1].view(B, 1, 1)) * 0.5
        fl = focal.reshape(-1, 1, 1)          # [B] or 0-dim (persp eval scripts)
        u = (u / fl).expand(B, H, W)
        v = (v / fl).expand(B, H, W)
        cam = torch.stack((u, -v, -torch.ones_like(u)), dim=-1)            # [B,H,W,3]
        prod = cam.unsqueeze(-2) * R.view(B, 1, 1, 3, 3)
        d = (prod[..., 0] + prod[..., 1]) + prod[..., 2]
        o = t.view(B, 1, 1, 3).expand(B, H, W, 3)
    else:                                     # nerf_utils.py:68-91
        u = (u - 0.5) * 2
        v = (v - 0.5) * 2
        if bbox is not None:
            u = bbox[:, 1, 0].view(B, 1, 1) * (u / 2 + 0.5) + bbox[:, 0, 0].view(B, 1, 1)
            v = -(bbox[:, 1, 1].view(B, 1, 1) * (-v / 2 + 0.5) + bbox[:, 0, 1].view(B, 1, 1))
        u = u.expand(B, H, W)
        v = v.expand(B, H, W)
        cam_o = torch.stack((u, -v, torch.zeros_like(u)), dim=-1)
        prod = cam_o.unsqueeze(-2) * R.view(B, 1, 1, 3, 3)
        o = ((prod[..., 0] + prod[..., 1]) + prod[..., 2]) + t.view(B, 1, 1, 3)
        zero = torch.zeros_like(u)
        cam_d = torch.stack((zero, zero, -torch.ones_like(u)), dim=-1)
        prod = cam_d.unsqueeze(-2) * R.view(B, 1, 1, 3, 3)
        d = ((prod[..., 0] + prod[..., 1]) + prod[..., 2]) / cam2world[:, 3, 3].view(B, 1, 1, 1)
    return o, d


def unit_directions(d, eps: float = 1e-12):
    """F.normalize(d, dim=-1)   (run.py:196)."""
    if d.requires_grad:
        return d / d.norm(p=2, dim=-1, keepdim=True).clamp_min(eps)
    return d / vec3_norm(d).clamp_min(eps).unsqueeze(-1)


# --------------------------------------------------------------------------------------
# A2  ray / scene-box near & far      lib/nerf_utils.py:227-275   (no grad, run.py:197)
# --------------------------------------------------------------------------------------
def box_near_far(o, d, scene_range: float, return_hit: bool = False):
    shape = o.shape[:-1]
    o = o.detach().reshape(-1, 3)
    d = d.detach().reshape(-1, 3)
    r = torch.tensor(scene_range, dtype=f32, device=o.device)
    inv = 1 / d
    neg = inv < 0
    lo = torch.where(neg, r, -r)              # plane met first along the ray
    hi = torch.where(neg, -r, r)
    tmin = (lo - o) * inv                     # [R,3]
    tmax = (hi - o) * inv
    hit = ~((tmin[:, 0] > tmax[:, 1]) | (tmin[:, 1] > tmax[:, 0]))
    near = torch.max(tmin[:, 0], tmin[:, 1])  # NaN-propagating, like the reference
    far = torch.min(tmax[:, 0], tmax[:, 1])
    hit = hit & ~((near > tmax[:, 2]) | (tmin[:, 2] > far))
    near = torch.max(near, tmin[:, 2])
    far = torch.min(far, tmax[:, 2])
    if not bool(hit.any()):
        # nerf_utils.py:260: `near_plane[mask].min()` on an empty selection raises.
        raise RuntimeError("compute_near_far_planes: no ray intersects the scene box")
    near = torch.where(hit, near, near[hit].min())
    far = torch.where(hit, far, far[hit].max())
    near = near.clamp(min=0.1)
    far = far.clamp(min=0.1)
    far = torch.where((far - near) < 1e-3, near + 1e-3, far)
    if return_hit:
        return near.reshape(shape), far.reshape(shape), hit.reshape(shape)
    return near.reshape(shape), far.reshape(shape)


# --------------------------------------------------------------------------------------
# A3  stratified depths + points      lib/nerf_utils.py:96-122
# --------------------------------------------------------------------------------------
def lerp_fma(a, b, w):
    """torch.lerp's CPU formula, restated (validated bitwise against torch.lerp)."""
    diff = (b - a).double()
    lo = (w.double() * diff + a.double()).float()
    hi = ((w - 1).double() * diff + b.double()).float()
    return torch.where(w < 0.5, lo, hi)


def stratified_depths(near, far, S: int, u: Optional[torch.Tensor] = None):
    """z[..., k] = lerp(near, far, k/S) (+ u * (far-near)/S when jittered)."""
    w = (torch.arange(S, dtype=torch.int64, device=near.device) / S).to(f32)
    z = lerp_fma(near.unsqueeze(-1), far.unsqueeze(-1), w)
    if u is not None:
        z = z + u * ((far - near).unsqueeze(-1) / S)
    return z


def points_on_rays(o, d, z):
    return o.unsqueeze(-2) + d.unsqueeze(-2) * z.unsqueeze(-1)


# --------------------------------------------------------------------------------------
# A4-A6  triplane field: gather, MLP, density + colour heads
#        models/generator.py:288-331, 587-681; models/stylegan.py:148-180
# --------------------------------------------------------------------------------------
@dataclass
class FieldParams:
    """Frozen description of the decoder + heads (weights are the *raw* EqualizedLinear tensors)."""
    W0: torch.Tensor              # [64, 32]   decoder.net[0].weight
    b0: torch.Tensor              # [64]
    W1: torch.Tensor              # [1+A or 1+3, 64]   decoder.net[2].weight
    b1: torch.Tensor
    scene_range: float
    use_sdf: bool = True
    attention_values: int = 10    # A; 0 -> plain wide-sigmoid RGB head
    beta: Optional[torch.Tensor] = None    # [1]
    alpha: Optional[torch.Tensor] = None   # [1]

    def gains(self):
        # stylegan.py:170-171 with lr_multiplier = 1
        return 1.0 / math.sqrt(self.W0.shape[1]), 1.0 / math.sqrt(self.W1.shape[1])


def triplane_features(planes, x):
    """planes [B,3,C,Hp,Wp] (xy, xz, yz); x [B,N,3] in box units -> mean feature [B,N,C].

    generator.py:312-328: bilinear, border padding, align_corners; plane `xy` is sampled
    at (x0 -> width, x1 -> height), `xz` at (x0, x2), `yz` at (x1, x2).
    """
    B, N, _ = x.shape
    g = x.view(B, N, 1, 3)
    acc = None
    for p, (a, b) in enumerate(((0, 1), (0, 2), (1, 2))):
        grid = torch.stack((g[..., a], g[..., b]), dim=-1)
        e = F.grid_sample(planes[:, p], grid, mode="bilinear", padding_mode="border",
                          align_corners=True)                      # [B,C,N,1]
        acc = e if acc is None else acc + e
    feat = acc / 3
    return feat.view(B, feat.shape[1], N).transpose(1, 2)


def decoder_mlp(feat, fp: FieldParams):
    g0, g1 = fp.gains()
    h = F.softplus(F.linear(feat, fp.W0 * g0, fp.b0))
    return F.linear(h, fp.W1 * g1, fp.b1)


def field_eval(planes, palette, pts, fp: FieldParams, want_probs: bool = False, want_normals: bool = False):
    """pts [B,N,3] world units -> sigma [B,N], rgb [B,N,3] (+ attention probs [B,N,A])
    (+ unit SDF gradients [B,N,3] when `want_normals`; everything is then detached, generator.py:599-623)."""
    if want_normals:                                                # generator.py:599-602
        assert fp.use_sdf
        pts = pts.detach().requires_grad_()
    x = pts / fp.scene_range                                        # generator.py:604
    with torch.no_grad():
        oob = (x.abs() > 1).any(dim=-1).to(f32)                     # generator.py:605-607
    y = decoder_mlp(triplane_features(planes, x), fp)
    dist, feats = y[..., 0], y[..., 1:]
    normals = None
    if want_normals:                                                # generator.py:614-623
        with torch.enable_grad():
            g, = torch.autograd.grad(dist.sum(), pts, create_graph=False)
        normals = F.normalize(g, dim=-1)
        dist, feats = dist.detach(), feats.detach()
    if fp.use_sdf:                                                  # generator.py:629-636
        t = -dist
        cdf = 0.5 + 0.5 * torch.sign(t) * (1 - torch.exp(-t.abs() / fp.beta))
        sigma = (1 / fp.alpha) * (cdf * (1 - oob))
    else:                                                           # generator.py:637-641
        sigma = F.softplus(dist - 1) * (1 - oob)
    probs = None
    if fp.attention_values > 0:                                     # generator.py:668-679
        probs = F.softmax(feats, dim=-1)
        rgb = torch.matmul(probs, palette)
    else:                                                           # generator.py:665-666
        rgb = torch.sigmoid(feats) * 2.004 - 1.002
    if want_normals:
        return sigma, rgb, probs, normals
    if want_probs:
        return sigma, rgb, probs
    return sigma, rgb


# --------------------------------------------------------------------------------------
# A7  transmittance weights + EG3D smoothing   nerf_utils.py:166-182, 20-25; run.py:266-272
# --------------------------------------------------------------------------------------
def segment_lengths(z, dnorm):
    """dist_i = (z[i+1]-z[i]) * |d|, last one = 0   (nerf_utils.py:137-142)."""
    gap = torch.cat((z[..., 1:] - z[..., :-1], torch.zeros_like(z[..., :1])), dim=-1)
    return gap * dnorm.unsqueeze(-1)


def transmittance_weights(sigma, z, dnorm):
    alpha = 1.0 - torch.exp(-sigma * segment_lengths(z, dnorm))
    keep = 1.0 - alpha + 1e-10
    if keep.requires_grad:
        T = torch.cumprod(keep[..., :-1], dim=-1)
    else:
        T = cumprod_f64(keep[..., :-1])
    T = torch.cat((torch.ones_like(T[..., :1]), T), dim=-1)
    return alpha * T


def smooth_weights(w):
    """max_pool1d(k=2,s=1,pad=1) -> avg_pool1d(k=2,s=1) -> +0.01   (run.py:266-272)."""
    ninf = torch.full_like(w[..., :1], float("-inf"))
    m = torch.maximum(torch.cat((ninf, w), dim=-1), torch.cat((w, ninf), dim=-1))   # S+1
    return (m[..., :-1] + m[..., 1:]) / 2 + 0.01


# --------------------------------------------------------------------------------------
# A8  inverse-CDF importance sampling          nerf_utils.py:185-224
# --------------------------------------------------------------------------------------
def inverse_cdf_samples(bins, weights, n: int, u: Optional[torch.Tensor] = None,
                        return_indices: bool = False):
    """bins [R,M], weights [R,M-1] -> n samples per row; `u=None` = deterministic linspace."""
    w = weights + 1e-5
    pdf = w / aten_row_sum(w).unsqueeze(-1)
    cdf = torch.cat((torch.zeros_like(pdf[..., :1]), cumsum_f64(pdf)), dim=-1)      # [R,M]
    if u is None:
        u = torch.linspace(0.0, 1.0, steps=n, dtype=w.dtype, device=w.device).expand(cdf.shape[0], n)
    u = u.contiguous()
    M = cdf.shape[-1]
    inds = (cdf.unsqueeze(1) <= u.unsqueeze(2)).sum(dim=-1)         # searchsorted(right=True)
    below = (inds - 1).clamp(min=0)
    above = inds.clamp(max=M - 1)
    c0, c1 = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    b0, b1 = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = c1 - c0
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - c0) / denom
    z = b0 + t * (b1 - b0)
    if return_indices:
        return z, inds, below, above
    return z


def importance_depths(sigma_c, z_c, dnorm, n: int, u: Optional[torch.Tensor] = None):
    """run.py:261-281 in one call: coarse sigma/z [R,S] -> fine depths [R,n] (no grad)."""
    with torch.no_grad():
        w = smooth_weights(transmittance_weights(sigma_c, z_c, dnorm))
        mid = 0.5 * (z_c[..., 1:] + z_c[..., :-1])
        return inverse_cdf_samples(mid, w[..., 1:-1], n, u)


# --------------------------------------------------------------------------------------
# A9  merge, A10 composite                     run.py:283-335; nerf_utils.py:125-163
# --------------------------------------------------------------------------------------
def merge_depths(z_c, z_f):
    return torch.sort(torch.cat((z_c, z_f), dim=-1), dim=-1, stable=True)


def composite(sigma, rgb, z, dnorm, white_background: bool, normals=None, semantics=None):
    w = transmittance_weights(sigma, z, dnorm)
    rgb_map = (w.unsqueeze(-1) * rgb).sum(dim=-2)
    depth_map = (w.detach() * z.detach()).sum(dim=-1)
    mask = w.sum(dim=-1)
    normal_map = None if normals is None else (w.detach().unsqueeze(-1) * normals).sum(dim=-2)
    sem_map = None if semantics is None else (w.unsqueeze(-1) * semantics).sum(dim=-2)
    if white_background:
        rgb_map = rgb_map + (1.0 - mask.unsqueeze(-1))
        if normal_map is not None:
            normal_map = normal_map + (1.0 - mask.unsqueeze(-1))
    return rgb_map, depth_map, mask, normal_map, sem_map


def zbuffer_depth(depth, d_unit, cam2world):
    """Ray-distance -> camera z depth (eval_kitti_persp.py:221-228, pose_utils.py:20-27)."""
    B = cam2world.shape[0]
    Rw2c = cam2world[:, :3, :3].transpose(-2, -1) / cam2world[:, 3:4, 3:4]
    prod = d_unit.unsqueeze(-2) * Rw2c.view(B, 1, 1, 3, 3)
    view = (prod[..., 0] + prod[..., 1]) + prod[..., 2]
    return (view * depth.unsqueeze(-1))[..., 2] * (-1)


# --------------------------------------------------------------------------------------
# N3  pose parametrisation -> cam2world          lib/pose_utils.py:30-75
# --------------------------------------------------------------------------------------
def pose_matrix(z0, t2, s, q, camera_flipped: bool):
    """(z0 [B] | None, t2 [B,2], s [B], unit quaternion q [B,4] = (w, x, y, z)) -> (cam2world [B,4,4],
    focal [B] | None).  Row i of the rotation block is the rotated basis vector e_i
    (pose_utils.py:30-46: v + 2 (w (u x v) + u x (u x v))); the translation column is R . t3 with
    t3 = (t2 / s, f / s), f = 1 + exp(z0) (perspective) or (t2, 1) / s (orthographic); a flipped camera
    negates columns 1..3 of the top three rows -- the translation included (pose_utils.py:60-62)."""
    B = q.shape[0]
    w, u = q[:, :1], q[:, 1:]
    eye = torch.eye(3, dtype=q.dtype, device=q.device).unsqueeze(0).expand(B, -1, -1)
    uu = u.unsqueeze(1).expand(-1, 3, -1)
    uv = torch.cross(uu, eye, dim=2)
    uuv = torch.cross(uu, uv, dim=2)
    R = eye + 2 * (w.unsqueeze(1) * uv + uuv)
    if z0 is not None:
        f = 1 + z0.exp()
        t3 = torch.cat((t2 / s.unsqueeze(-1), (f / s).unsqueeze(-1)), dim=-1)
        focal = f / 2
    else:
        # the reference divides [B,3] by s [B] WITHOUT unsqueezing (pose_utils.py:69): it broadcasts over
        # the last dimension and is only the intended per-image division for B = 1 (it raises for
        # B not in {1, 3}); restated as written
        t3 = torch.cat((t2, torch.ones_like(t2[:, :1])), dim=-1) / s
        focal = None
    trans = (t3[:, None, :] * R).sum(dim=-1)
    top = torch.cat((R, trans.unsqueeze(-1)), dim=-1)                  # [B,3,4]
    if camera_flipped:
        top = torch.cat((top[..., :1], -top[..., 1:]), dim=-1)
    bottom = torch.tensor([0.0, 0.0, 0.0, 1.0], dtype=q.dtype, device=q.device).expand(B, 1, 4)
    return torch.cat((top, bottom), dim=1), focal


# --------------------------------------------------------------------------------------
# A11  the whole render                        run.py:176-350
# --------------------------------------------------------------------------------------
def render(planes, palette, fp: FieldParams, H: int, W: int, cam2world, focal=None, center=None,
           bbox=None, S: int = 64, white_background: bool = False, u_strat=None, u_pdf=None,
           randomize: bool = True, fine_sampling: bool = True, no_cam_grad: bool = False,
           zbuffer: bool = False, want_semantics: bool = False, return_aux: bool = False,
           want_normals: bool = False, want_coords: bool = False):
    """Full hot path.  `randomize=True` needs explicit uniforms u_strat [B,H,W,S], u_pdf [R,S].
    `want_normals` / `want_semantics` / `want_coords` = compute_normals / compute_semantics /
    compute_coords of run.py:186-188; the maps come back in aux["normals"], aux["semantics"] (coords
    take the semantics slot, run.py:337-338)."""
    B = cam2world.shape[0]
    R = B * H * W
    o, d_raw = pixel_rays(H, W, cam2world, focal, center, bbox)
    d = unit_directions(d_raw)
    near, far = box_near_far(o, d, fp.scene_range)
    if randomize:
        assert u_strat is not None
    z_c = stratified_depths(near, far, S, u_strat if randomize else None)           # [B,H,W,S]
    pts = points_on_rays(o, d, z_c)
    if no_cam_grad:                                                  # run.py:211-214
        pts, d = pts.detach(), d.detach()
    dn = d.norm(p=2, dim=-1) if d.requires_grad else vec3_norm(d)
    out = field_eval(planes, palette, pts.reshape(B, H * W * S, 3), fp, want_probs=True,
                     want_normals=want_normals)
    sigma = out[0].view(B, H, W, S)
    rgb = out[1].view(B, H, W, S, 3)
    sem = out[2].view(B, H, W, S, -1) if (want_semantics and out[2] is not None) else None
    nrm = out[3].view(B, H, W, S, 3) if want_normals else None
    crd = pts.reshape(B, H, W, S, 3) if want_coords else None       # generator.py:644
    aux = {"origins": o, "dirs": d, "near": near, "far": far, "z_coarse": z_c}
    z_all = z_c
    if fine_sampling:
        z_f = importance_depths(sigma.detach().reshape(R, S), z_c.reshape(R, S),
                                dn.detach().reshape(R), S,
                                (u_pdf if randomize else None)).view(B, H, W, S)
        z_all, perm = merge_depths(z_c, z_f)
        pts_f = points_on_rays(o, d, z_f) if not no_cam_grad else points_on_rays(o, d, z_f).detach()
        out_f = field_eval(planes, palette, pts_f.reshape(B, H * W * S, 3), fp, want_probs=True,
                           want_normals=want_normals)
        sigma = torch.cat((sigma, out_f[0].view(B, H, W, S)), dim=-1).gather(-1, perm)
        rgb = torch.cat((rgb, out_f[1].view(B, H, W, S, 3)), dim=-2).gather(
            -2, perm.unsqueeze(-1).expand(-1, -1, -1, -1, 3))
        if sem is not None:
            sem = torch.cat((sem, out_f[2].view(B, H, W, S, -1)), dim=-2).gather(
                -2, perm.unsqueeze(-1).expand(-1, -1, -1, -1, sem.shape[-1]))
        p3 = perm.unsqueeze(-1).expand(-1, -1, -1, -1, 3)
        if nrm is not None:
            nrm = torch.cat((nrm, out_f[3].view(B, H, W, S, 3)), dim=-2).gather(-2, p3)
        if crd is not None:
            crd = torch.cat((crd, pts_f.reshape(B, H, W, S, 3)), dim=-2).gather(-2, p3)
        aux.update(z_fine=z_f, perm=perm)
    if crd is not None:                                              # run.py:337-338
        sem = crd
    rgb_map, depth, mask, nrm_map, sem_map = composite(sigma, rgb, z_all, dn, white_background,
                                                       normals=nrm, semantics=sem)
    aux["normals"] = nrm_map
    if zbuffer:
        depth = zbuffer_depth(depth, d, cam2world)
    aux.update(z_sorted=z_all, semantics=sem_map)
    if return_aux:
        return rgb_map, depth, mask, aux
    return rgb_map, depth, mask
