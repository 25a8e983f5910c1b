"""Helpers shared by the golden-vector tests (CPU oracle tests and GPU parity tests)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def eq_bits(a, b):
    """Bitwise-equal values, NaN == NaN."""
    return a.shape == b.shape and bool(((a == b) | (a.isnan() & b.isnan())).all())


def rel_err(a, b):
    """max |a-b| / max |b|  -- the 'relative' error the 1e-4 (fp32) / 1e-2 (bf16) contract uses."""
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


RAY_VARIANTS = {
    # name: (camera key, perspective, bbox, center)
    "persp": ("cam_persp", True, False, False), "persp_c": ("cam_persp", True, False, True),
    "persp_b": ("cam_persp", True, True, False), "persp_bc": ("cam_persp", True, True, True),
    "ortho": ("cam_ortho", False, False, False), "ortho_b": ("cam_ortho", False, True, False),
    "cub": ("cam_cub", False, False, False),
}
NEARFAR_RANGE = {"persp": 1.0, "ortho": 0.6, "cub": 2.0}


def render_case(name):
    g = load("render_" + name)
    H, W, S, persp, randomize, white, no_cam, zbuf = [int(v) for v in g["meta"]]
    g["cfg"] = dict(H=H, W=W, S=S, persp=bool(persp), randomize=bool(randomize), white=bool(white),
                    no_cam=bool(no_cam), zbuffer=bool(zbuf), scene_range=float(g["scene_range"]))
    return g


RENDER_CASES = ["persp_rand", "ortho_det", "persp_white_nocam", "kitti_zbuf", "persp_tiles", "shapenet_tiles", "cub_tiles"]
# forward-only fixtures of the extra render outputs; g["flags"] = (normals, semantics, coords)
N1_CASES = ["persp_n1", "ortho_coords"]
