"""Drop-in for the reference's `lib/nerf_utils.py`: the same seven function names, argument
order and return values (lib/nerf_utils.py:20-275), each backed by sm_100a kernels through the C ABI.

Install with `nfi_b200.patch.install()` (puts this module at `sys.modules['lib.nerf_utils']`) or
import it directly.  CUDA tensors only -- there is no CPU path here.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops


def cumprod_exclusive(tensor: torch.Tensor) -> torch.Tensor:
    """lib/nerf_utils.py:20-25: exclusive product scan along the last dim (forward only -- the
    reference only calls it inside render_volume_density*, which are replaced wholesale below)."""
    return ops.cumprod_exclusive(tensor)


def get_ray_bundle(height: int, width: int, focal_length: Optional[torch.Tensor],
                   tform_cam2world: torch.Tensor, bbox: Optional[torch.Tensor],
                   center: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """lib/nerf_utils.py:28-93 -> (ray_origins, ray_directions) [B,H,W,3] (directions NOT normalised)."""
    return ops.ray_bundle(height, width, focal_length, tform_cam2world, bbox, center, normalize=False)


def compute_query_points_from_rays(ray_origins: torch.Tensor, ray_directions: torch.Tensor,
                                   near_thresh: torch.Tensor, far_thresh: torch.Tensor,
                                   num_samples: int, randomize: bool = True
                                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """lib/nerf_utils.py:96-122 -> (query_points [..,S,3], depth_values [..,S]).

    The uniforms are drawn exactly like the reference (`torch.rand_like(depth_values)`), on the
    tensors' device, so a seeded reference run on the same device sees the same numbers."""
    u = None
    if randomize:
        u = torch.rand((*near_thresh.shape, num_samples), dtype=torch.float32, device=near_thresh.device)
    z = ops.stratified(near_thresh, far_thresh, num_samples, randomize, u)
    pts = ops.points_on_rays(ray_origins, ray_directions, z)
    return pts, z


def render_volume_density(sigma_a: torch.Tensor, rgb: torch.Tensor, ray_origins: torch.Tensor,
                          ray_directions: torch.Tensor, depth_values: torch.Tensor,
                          normals: Optional[torch.Tensor] = None,
                          semantics: Optional[torch.Tensor] = None, white_background: bool = True):
    """lib/nerf_utils.py:125-163 -> (rgb_map, depth_map, mask, normal_map|None, semantic_map|None)."""
    return ops.composite(sigma_a, rgb, ray_directions, depth_values, normals, semantics,
                         white_background)


def render_volume_density_weights_only(sigma_a: torch.Tensor, ray_origins: torch.Tensor,
                                       ray_directions: torch.Tensor,
                                       depth_values: torch.Tensor) -> torch.Tensor:
    """lib/nerf_utils.py:166-182 (called under no_grad by render, run.py:261-265)."""
    return ops.weights_only(sigma_a, ray_directions, depth_values, smooth=False)


def sample_pdf(bins: torch.Tensor, weights: torch.Tensor, num_samples: int,
               deterministic: bool = False) -> torch.Tensor:
    """lib/nerf_utils.py:185-224.  `torch.rand([R, num_samples])` as in the reference (:204)."""
    u = None
    if not deterministic:
        u = torch.rand((bins.shape[0], num_samples), dtype=torch.float32, device=bins.device)
    return ops.sample_pdf(bins, weights, num_samples, u)


def compute_near_far_planes(ray_origins: torch.Tensor, ray_directions: torch.Tensor,
                            scene_range: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """lib/nerf_utils.py:227-275; raises RuntimeError when no ray hits the box (as :260 does)."""
    return ops.near_far(ray_origins, ray_directions, scene_range, check_hits=True)
