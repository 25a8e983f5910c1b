"""CPU emulation of the kernels' __host__ __device__ arithmetic (tests/hostcheck/hostcheck.cu) against
the reference goldens / the oracle.  Test infrastructure: proves the math the CUDA kernels execute
(ray generation, slab test, lerp, bilinear footprint, MLP, heads and all adjoint formulas) before any
GPU time is spent.  Needs nvcc (host compile only); skipped where it is absent."""
import ctypes
import os
import shutil
import subprocess

import numpy as np
import pytest
import torch

import nfi_oracle as O
from golden_io import NEARFAR_RANGE, RAY_VARIANTS, eq_bits, load, rel_err

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostcheck", "hostcheck.cu")
LIB = os.path.join(HERE, "hostcheck", "libhostcheck.so")
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


@pytest.fixture(scope="module")
def hc():
    if not os.path.exists(NVCC):
        pytest.skip("nvcc not available")
    csrc = os.path.join(os.path.dirname(HERE), "nerf-from-image_b200", "csrc")
    deps = [SRC] + [os.path.join(csrc, f) for f in ("nfi_math.cuh", "nfi_field_math.cuh")]
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        subprocess.check_call([NVCC, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-Xcompiler",
                               "-fPIC,-ffp-contract=off", "-shared", "-o", LIB, SRC])
    return ctypes.CDLL(LIB)


def P(t):
    if t is None:
        return None
    assert t.dtype == torch.float32 and t.is_contiguous()
    return ctypes.c_void_p(t.data_ptr())


@pytest.fixture(scope="module")
def fn():
    return load("functions")


@pytest.mark.parametrize("variant", sorted(RAY_VARIANTS))
def test_rays_bitexact(hc, fn, variant):
    cam_key, persp, bb, c = RAY_VARIANTS[variant]
    H, W = [int(v) for v in fn["HW"]]
    cam = fn[cam_key].contiguous()
    B = cam.shape[0]
    o = torch.empty(B * H * W, 3); d = torch.empty(B * H * W, 3)
    hc.hc_rays(P(cam), P(fn["focal"]) if persp else None, 1, P(fn["center"]) if c else None,
               P(fn["bbox"].contiguous()) if bb else None, B, H, W, 0, P(o), P(d))
    assert eq_bits(o.view(B, H, W, 3), fn["rays_o_" + variant])
    assert eq_bits(d.view(B, H, W, 3), fn["rays_d_" + variant])


@pytest.mark.parametrize("variant", sorted(NEARFAR_RANGE))
def test_near_far_stratified_bitexact(hc, fn, variant):
    cam_key, persp, _, _ = RAY_VARIANTS[variant]
    H, W = [int(v) for v in fn["HW"]]
    cam = fn[cam_key].contiguous()
    B = cam.shape[0]; R = B * H * W
    o = torch.empty(R, 3); d = torch.empty(R, 3)
    hc.hc_rays(P(cam), P(fn["focal"]) if persp else None, 1, None, None, B, H, W, 1, P(o), P(d))
    assert eq_bits(d.view(B, H, W, 3), fn["unit_d_" + variant])
    near = torch.empty(R); far = torch.empty(R)
    hc.hc_near_far(P(o), P(d), ctypes.c_int64(R), ctypes.c_float(NEARFAR_RANGE[variant]), P(near), P(far))
    assert eq_bits(near.view(B, H, W), fn["near_" + variant]) and eq_bits(far.view(B, H, W), fn["far_" + variant])
    if variant == "persp":
        z = torch.empty(R, 64)
        hc.hc_stratified(P(near), P(far), P(fn["strat_u"].reshape(R, 64).contiguous()), ctypes.c_int64(R), 64, P(z))
        assert eq_bits(z.view(B, H, W, 64), fn["strat_z_rand"])
        pts = torch.empty(R, 64, 3)
        hc.hc_points(P(o), P(d), P(z), ctypes.c_int64(R), 64, P(pts))
        assert eq_bits(pts.view(B, H, W, 64, 3), fn["strat_pts_rand"])
        hc.hc_stratified(P(near), P(far), None, ctypes.c_int64(R), 64, P(z))
        assert eq_bits(z.view(B, H, W, 64), fn["strat_z_det"])


@pytest.mark.parametrize("kind", ["persp", "persp_bc", "ortho", "cub"])
def test_rays_backward(hc, fn, kind):
    cam_key, persp, bb, c = RAY_VARIANTS[kind]
    H, W = 6, 8
    cam = fn[cam_key].clone().requires_grad_()
    focal = fn["focal"].clone().requires_grad_() if persp else None
    B = cam.shape[0]
    o, d = O.pixel_rays(H, W, cam, focal, fn["center"] if c else None, fn["bbox"] if bb else None)
    dn = O.unit_directions(d)
    g = torch.Generator().manual_seed(3)
    go = torch.randn(B * H * W, 3, generator=g); gd = torch.randn(B * H * W, 3, generator=g)
    wants = [cam] + ([focal] if persp else [])
    ref = torch.autograd.grad((o.reshape(-1, 3) * go).sum() + (dn.reshape(-1, 3) * gd).sum(), wants)
    gcam = torch.empty(B, 16); gf = torch.empty(B)
    hc.hc_rays_bwd(P(cam.detach().contiguous()), P(focal.detach()) if persp else None, 1,
                   P(fn["center"]) if c else None, P(fn["bbox"].contiguous()) if bb else None, B, H, W, 1,
                   P(go), P(gd), P(gcam), P(gf))
    got = gcam.view(B, 4, 4).clone()
    got[:, 3, :3] = 0      # bottom row is not an input of the ray generator
    assert rel_err(got, ref[0]) < 1e-5
    if persp:
        assert rel_err(gf, ref[1]) < 1e-5


def to_cl(planes):      # [B,96,H,W] -> [B,3,H,W,32]
    B, _, H, W = planes.shape
    return planes.view(B, 3, 32, H, W).permute(0, 1, 3, 4, 2).contiguous()


@pytest.mark.parametrize("name,A,use_sdf", [("sdf_attn", 10, True), ("density_rgb", 0, False)])
def test_field_forward_backward(hc, name, A, use_sdf):
    g = load("decoder_" + name)
    B, res = g["planes"].shape[0], g["planes"].shape[-1]
    N = g["pts"].shape[1]
    n_out = g["W1"].shape[0]
    pcl = to_cl(g["planes"])
    common = [P(g["W0"]), P(g["b0"]), P(g["W1"]), P(g["b1"]), ctypes.c_float(float(g["beta"])),
              ctypes.c_float(float(g["alpha"])), n_out, A, int(use_sdf)]
    # raw decoder (mode 1): coords already in box units -> scene_range 1
    coords = (g["pts"] / 1.4).contiguous()
    y = torch.empty(B, N, n_out)
    gpl = torch.zeros_like(pcl); gpt = torch.empty(B, N, 3)
    hc.hc_field(*common, ctypes.c_float(1.0), P(pcl), None, B, res, res, P(coords), ctypes.c_int64(N), 1, P(y),
                None, P(g["gy"].contiguous()), P(gpl), None, P(gpt), None)
    assert rel_err(y, g["y"]) < 1e-5
    assert rel_err(gpl, to_cl(g["dy_planes"])) < 1e-5
    assert rel_err(gpt / 1.4, g["dy_pts"]) < 1e-5
    # heads (mode 0)
    out = torch.empty(B, N, 4); probs = torch.empty(B, N, max(A, 1))
    pal = g["palette"].contiguous() if A > 0 else None
    gout = torch.cat((g["gsigma"].unsqueeze(-1), g["grgb"]), dim=-1).contiguous()
    gpl = torch.zeros_like(pcl); gpal = torch.zeros(B, max(A, 1), 3); gba = torch.zeros(2)
    hc.hc_field(*common, ctypes.c_float(1.4), P(pcl), P(pal), B, res, res, P(g["pts"].contiguous()),
                ctypes.c_int64(N), 0, P(out), P(probs) if A > 0 else None, P(gout), P(gpl),
                P(gpal) if A > 0 else None, P(gpt), P(gba))
    assert rel_err(out[..., 0], g["sigma"]) < 1e-5 and rel_err(out[..., 1:], g["rgb"]) < 1e-5
    assert eq_bits((out[..., 0] == 0), (g["sigma"] == 0))          # out-of-box mask decisions
    assert rel_err(gpl, to_cl(g["ds_planes"])) < 2e-5
    assert rel_err(gpt, g["ds_pts"]) < 2e-5
    if A > 0:
        assert rel_err(probs, g["semantics"]) < 1e-5
        assert rel_err(gpal, g["ds_palette"]) < 2e-5
    if use_sdf:
        assert abs(float(gba[0]) - float(g["ds_beta"])) < 2e-4 * abs(float(g["ds_beta"]))
        assert abs(float(gba[1]) - float(g["ds_alpha"])) < 2e-4 * abs(float(g["ds_alpha"]))
