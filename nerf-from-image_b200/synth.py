"""Synthetic inputs of the shapes BASELINE.json names.

Data generation only -- nothing here is on the render path.  There is no network
for datasets or checkpoints, so tests, `bench.py` and the golden-vector script all
draw their inputs from these seeded generators:

  * planes: smooth bicubic-upsampled Gaussian noise, shaped like the StyleGAN2
    synthesis output `[B, 96, res, res]` (models/generator.py:475-477);
  * decoder: random-init EqualizedLinear weights (models/stylegan.py:161-167) with the
    SDF bias shifted so that ~35 % of the box is inside the surface -- a random-init
    generator renders an almost empty image, which would make parity vacuous;
  * palette: attention values in [-1.002, 1.002] (models/generator.py:183-184);
  * cameras: the optimisable pose parametrisation (z0, t2, s, q) of
    lib/pose_utils.py:48-75, restated, so pose gradients flow through the real path.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return g


def make_planes(B: int, res: int = 256, seed: int = 1, channels: int = 96, coarse: int = 16):
    base = torch.randn(B, channels, coarse, coarse, generator=_gen(seed))
    return F.interpolate(base, size=(res, res), mode="bicubic", align_corners=True).contiguous()


def _field_sdf_torch(planes, W0, b0, W1, b1, pts_unit):
    """Plain-torch evaluation of decoder output 0 -- used ONLY to calibrate the bias below."""
    B, N, _ = pts_unit.shape
    g = pts_unit.view(B, N, 1, 3)
    p5 = planes.view(B, 3, planes.shape[1] // 3, planes.shape[-2], planes.shape[-1])
    acc = 0
    for p, (a, b) in enumerate(((0, 1), (0, 2), (1, 2))):
        acc = acc + F.grid_sample(p5[:, p], torch.stack((g[..., a], g[..., b]), -1),
                                  mode="bilinear", padding_mode="border", align_corners=True)
    feat = (acc / 3).view(B, -1, N).transpose(1, 2)
    h = F.softplus(F.linear(feat, W0 / math.sqrt(W0.shape[1]), b0))
    return F.linear(h, W1 / math.sqrt(W1.shape[1]), b1)[..., 0]


def make_field(planes: Optional[torch.Tensor], seed: int = 3, attention_values: int = 10,
               feat_ch: int = 32, hidden: int = 64, inside_fraction: float = 0.35,
               use_sdf: bool = True):
    """Decoder weights (raw EqualizedLinear tensors) + beta/alpha, SDF bias calibrated on `planes`."""
    g = _gen(seed)
    n_out = 1 + (attention_values if attention_values > 0 else 3)
    W0 = torch.randn(hidden, feat_ch, generator=g)
    b0 = torch.randn(hidden, generator=g) * 0.1
    W1 = torch.randn(n_out, hidden, generator=g)
    b1 = torch.randn(n_out, generator=g) * 0.1
    if planes is not None:
        pts = torch.rand(1, 20000, 3, generator=g) * 2 - 1
        sdf = _field_sdf_torch(planes[:1], W0, b0, W1, b1, pts)
        if use_sdf:      # inside <=> sdf < 0: move the `inside_fraction` quantile to 0
            b1[0] -= torch.quantile(sdf.flatten(), inside_fraction)
        else:            # density head is softplus(d - 1): put the same fraction above d = 3
            b1[0] -= torch.quantile(sdf.flatten(), 1 - inside_fraction) - 3.0
    return {"W0": W0, "b0": b0, "W1": W1, "b1": b1,
            "beta": torch.tensor([0.1]), "alpha": torch.tensor([1.0])}


def make_palette(B: int, attention_values: int = 10, seed: int = 4):
    return (torch.rand(B, attention_values, 3, generator=_gen(seed)) * 2 - 1) * 1.002


# ---- caller-side pose parametrisation (lib/pose_utils.py:28-75, restated) -------------------
def quaternion_to_matrix(q):
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rows = [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]
    # the reference rotates the rows of eye(3) by q, i.e. it builds R^T of the usual convention
    return torch.stack(rows, dim=-1).view(-1, 3, 3).transpose(1, 2)


def pose_to_matrix(z0, t2, s, q, camera_flipped: bool):
    """(z0, t2, s, q) -> cam2world [B,4,4], focal [B] (None for the orthographic model)."""
    R = quaternion_to_matrix(q)
    B = q.shape[0]
    if z0 is not None:
        f = 1 + z0.exp()
        t3 = torch.cat((t2 / s.unsqueeze(-1), (f / s).unsqueeze(-1)), dim=-1)
        focal = f / 2
    else:
        t3 = torch.cat((t2, torch.ones_like(t2[:, :1])), dim=-1) / s.unsqueeze(-1)
        focal = None
    trans = (t3[:, None, :] * R).sum(dim=-1)
    top = torch.cat((R, trans.unsqueeze(-1)), dim=-1)
    if camera_flipped:
        sign = torch.tensor([1.0, -1.0, -1.0, -1.0], dtype=top.dtype, device=top.device)
        top = top * sign
    bottom = torch.tensor([0.0, 0.0, 0.0, 1.0], dtype=top.dtype, device=top.device).expand(B, 1, 4)
    return torch.cat((top, bottom), dim=1), focal


def make_pose_params(B: int, seed: int = 2, perspective: bool = True,
                     focal_range=(0.9, 2.3), dist_jitter=(0.8, 1.2), ortho_scale=(0.5, 1.0)):
    """Random object-centred poses: unit quaternion, small in-plane offset, camera looking at 0."""
    g = _gen(seed)
    q = F.normalize(torch.randn(B, 4, generator=g), dim=-1)
    t2 = (torch.rand(B, 2, generator=g) - 0.5) * 0.1
    if perspective:
        focal = focal_range[0] + (focal_range[1] - focal_range[0]) * torch.rand(B, generator=g)
        z0 = torch.log(2 * focal - 1)
        s = 1.0 / (dist_jitter[0] + (dist_jitter[1] - dist_jitter[0]) * torch.rand(B, generator=g))
    else:
        z0 = None
        s = ortho_scale[0] + (ortho_scale[1] - ortho_scale[0]) * torch.rand(B, generator=g)
    return z0, t2, s, q


def make_cameras(B: int, seed: int = 2, perspective: bool = True, camera_flipped: bool = True,
                 **kw):
    z0, t2, s, q = make_pose_params(B, seed, perspective, **kw)
    cam, focal = pose_to_matrix(z0, t2, s, q, camera_flipped)
    return cam, focal


def make_ortho_cub_cameras(B: int, seed: int = 2):
    """CUB-style orthographic cam2world with T[3,3] = 1/scale and a z offset of 10
    (data/datasets.py:556-567, restated): world2cam = flip @ [s*Q | s*(t + 10 e_z); 0 | s],
    inverted in closed form.  Exercises the `/ T[3,3]` branch of the ray generator."""
    g = _gen(seed)
    Q = quaternion_to_matrix(F.normalize(torch.randn(B, 4, generator=g), dim=-1))
    s = 0.5 + 0.5 * torch.rand(B, generator=g)
    t = torch.cat(((torch.rand(B, 2, generator=g) - 0.5) * 0.2, torch.full((B, 1), 10.0)), dim=-1)
    flip = torch.diag(torch.tensor([1.0, -1.0, -1.0]))
    cam = torch.zeros(B, 4, 4)
    cam[:, :3, :3] = (Q.transpose(1, 2) @ flip) / s.view(B, 1, 1)
    cam[:, :3, 3] = -(Q.transpose(1, 2) @ t.unsqueeze(-1)).squeeze(-1) / s.view(B, 1)
    cam[:, 3, 3] = 1.0 / s
    return cam
