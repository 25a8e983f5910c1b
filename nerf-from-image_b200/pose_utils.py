"""Drop-in for the part of `lib/pose_utils.py` that sits on the inversion step:
`pose_to_matrix` (lib/pose_utils.py:48-75), optionally with the `F.normalize(R_, dim=-1)` the loop
applies to the quaternion first (run.py:2257-2262) folded in -- one kernel forward, one backward,
instead of ~30 small ATen launches per step.  Same name, arguments and return tuple."""
from __future__ import annotations

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from ._lib import call, f32c, ptr


class _PoseToMatrix(Function):
    @staticmethod
    def forward(ctx, z0, t2, s, q, camera_flipped, normalize_q):
        t2c, sc, qc = f32c(t2), f32c(s), f32c(q)
        z0c = f32c(z0)
        B = qc.shape[0]
        cam = torch.empty(B, 4, 4, dtype=torch.float32, device=qc.device)
        focal = torch.empty(B, dtype=torch.float32, device=qc.device) if z0 is not None else None
        call("nfi_pose_to_matrix_fwd", qc, ptr(z0c), ptr(t2c), ptr(sc), ptr(qc), B, int(bool(camera_flipped)),
             int(bool(normalize_q)), ptr(cam), ptr(focal))
        ctx.save_for_backward(z0c, t2c, sc, qc)
        ctx.flags = (int(bool(camera_flipped)), int(bool(normalize_q)))
        return cam, focal

    @staticmethod
    @once_differentiable
    def backward(ctx, g_cam, g_focal):
        z0, t2, s, q = ctx.saved_tensors
        B = q.shape[0]
        g_cam = f32c(g_cam) if g_cam is not None else torch.zeros(B, 4, 4, dtype=torch.float32, device=q.device)
        g_focal = f32c(g_focal)
        g_z0 = torch.empty_like(z0) if z0 is not None else None
        g_t2, g_s, g_q = torch.empty_like(t2), torch.empty_like(s), torch.empty_like(q)
        call("nfi_pose_to_matrix_bwd", q, ptr(z0), ptr(t2), ptr(s), ptr(q), B, ctx.flags[0], ctx.flags[1],
             ptr(g_cam), ptr(g_focal), ptr(g_z0), ptr(g_t2), ptr(g_s), ptr(g_q))
        return g_z0, g_t2, g_s, g_q, None, None


def pose_to_matrix(z0, t2, s, q, camera_flipped: bool, normalize_quaternion: bool = False):
    """(z0 [B] | None, t2 [B,2], s [B], q [B,4], camera_flipped) -> (cam2world [B,4,4], focal [B] | None),
    lib/pose_utils.py:48-75.  `normalize_quaternion=True` = pose_to_matrix(..., F.normalize(q, dim=-1), ...)."""
    if z0 is None and q.shape[0] != 1:
        # pose_utils.py:69 divides [B,3] by s [B] without unsqueezing: an error for B not in {1, 3}
        raise RuntimeError("pose_to_matrix: the orthographic branch of the reference is only defined for one "
                           "image per call (lib/pose_utils.py:69); got B = %d" % q.shape[0])
    return _PoseToMatrix.apply(z0, t2, s, q, camera_flipped, normalize_quaternion)
