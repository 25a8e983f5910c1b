"""The fused C-ABI render (`nfi_render_fwd` / `nfi_render_bwd`, include/nfi_b200.h): run.py::render
(run.py:176-350) and its backward as one call each, inside one workspace tensor.

This is the entry point a non-Python host binds; from Python it serves callers that do not need
autograd -- `torch.no_grad()` rendering (FID / visualisation, run.py:1121) -- and hosts that drive the
backward themselves.  Same kernels, same results as `render.render_planes`.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib, ops
from ._lib import f32c, ptr
from .decoder import Field


def _vp(t):
    p = ptr(t)
    return None if p is None else p.value


class CRender:
    def __init__(self, planes_cl, palette, field: Field, height, width, tform_cam2world, focal_length=None,
                 center=None, bbox=None, depth_samples_per_ray=64, randomize=True, white_background=False,
                 zbuffer_depth=False, u_strat=None, u_pdf=None, seed=0, offset=0, with_backward=True):
        self.pl = f32c(planes_cl.detach())
        B, three, Hp, Wp, C = self.pl.shape
        assert three == 3 and C == 32
        self.B, self.H, self.W, self.S, self.Hp, self.Wp = B, int(height), int(width), int(depth_samples_per_ray), Hp, Wp
        self.dev = self.pl.device
        cam = f32c(tform_cam2world.detach())
        assert cam.shape == (B, 4, 4)
        focal = None if focal_length is None else f32c(focal_length.detach()).reshape(-1)
        self.focal_stride = 0 if (focal is not None and focal.numel() == 1 and B > 1) else 1
        wts = [t.detach() if t is not None else None for t in field.tensors()]
        fstruct, keep = _lib.make_field_struct(*wts, field.cfg.attention_values, field.cfg.use_sdf,
                                               field.cfg.scene_range)
        self._keep = [keep, cam, focal, f32c(palette.detach()) if palette is not None else None,
                      f32c(center.detach()) if center is not None else None,
                      f32c(bbox.detach()) if bbox is not None else None,
                      f32c(u_strat) if u_strat is not None else None, f32c(u_pdf) if u_pdf is not None else None]
        _, cam, focal, pal, cen, bb, us, up = self._keep
        self.n_out, self.A = fstruct.n_out, field.cfg.attention_values
        self.p = _lib.NfiRender(fstruct, _vp(self.pl), _vp(pal), _vp(cam), _vp(focal), _vp(cen), _vp(bb), _vp(us),
                                _vp(up), int(seed), int(offset), B, self.H, self.W, self.S, Hp, Wp,
                                self.focal_stride, int(bool(randomize)), int(bool(white_background)),
                                int(bool(zbuffer_depth)))
        l = _lib.lib()
        nbytes = int(l.nfi_render_workspace_bytes(ctypes.byref(self.p), int(bool(with_backward))))
        if nbytes == 0:
            raise _lib.NfiError("nfi_render_workspace_bytes: malformed parameters")
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)     # torch allocations are 512-byte aligned
        self.with_backward = bool(with_backward)

    def forward(self, check_hits=True):
        B, H, W = self.B, self.H, self.W
        rgb = torch.empty(B, H, W, 3, dtype=torch.float32, device=self.dev)
        depth = torch.empty(B, H, W, dtype=torch.float32, device=self.dev)
        mask = torch.empty(B, H, W, dtype=torch.float32, device=self.dev)
        _lib.call("nfi_render_fwd", self.pl, ctypes.byref(self.p), ptr(self.ws), ptr(rgb), ptr(depth), ptr(mask))
        if check_hits and int(self.ws[:4].view(torch.int32).item()) == 0:
            raise RuntimeError(ops._NO_HIT_MSG)          # lib/nerf_utils.py:260
        return rgb, depth, mask

    def backward(self, g_rgb, g_mask=None, want_planes=True, want_palette=True, want_pose=True,
                 want_weights=False, want_params=False):
        """-> dict(planes_cl, palette, cam2world, focal, params) of the requested gradients."""
        assert self.with_backward
        B = self.B
        g_rgb = f32c(g_rgb)
        g_mask = f32c(g_mask)
        out = {}
        out["planes_cl"] = torch.zeros_like(self.pl) if want_planes else None
        out["palette"] = torch.zeros(B, self.A, 3, dtype=torch.float32, device=self.dev) \
            if (want_palette and self.A > 0) else None
        out["cam2world"] = torch.empty(B, 4, 4, dtype=torch.float32, device=self.dev) if want_pose else None
        out["focal"] = torch.empty(B, dtype=torch.float32, device=self.dev) if want_pose else None
        out["params"] = torch.zeros(sum(ops._param_sizes(self.n_out)), dtype=torch.float32, device=self.dev) \
            if (want_weights or want_params) else None
        _lib.call("nfi_render_bwd", self.pl, ctypes.byref(self.p), ptr(self.ws), ptr(g_rgb), ptr(g_mask),
                  ptr(out["planes_cl"]), ptr(out["palette"]), ptr(out["cam2world"]), ptr(out["focal"]),
                  ptr(out["params"]), int(bool(want_weights)))
        return out
