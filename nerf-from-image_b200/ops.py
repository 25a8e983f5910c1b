"""torch.autograd.Function wrappers, one per kernel group of libnfi_b200.so.

Each forward/backward is one (or a few) C-ABI calls on the caller's current CUDA stream; every
buffer -- outputs, workspaces, saved-for-backward state -- is a torch tensor, so the caching
allocator and autograd own all lifetimes ("Ownership").  Backward is
once-differentiable.  Nothing here computes on the CPU or falls back to PyTorch ops.
"""
from __future__ import annotations

import threading

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import _lib
from ._lib import call, f32c, ptr

NEARFAR_WS_BYTES = 16384


def _empty(ref, *shape, dtype=torch.float32):
    return torch.empty(shape, dtype=dtype, device=ref.device)


# ---------------------------------------------------------------------------------------------
# A1  rays     lib/nerf_utils.py:28-93 (+ F.normalize, run.py:196)
# ---------------------------------------------------------------------------------------------
class _RayBundle(Function):
    @staticmethod
    def forward(ctx, cam2world, focal, center, bbox, H, W, normalize):
        cam = f32c(cam2world)
        B = cam.shape[0]
        fo = None if focal is None else f32c(focal.reshape(-1))
        fstride = 0
        if fo is not None:
            assert fo.numel() in (1, B), "focal_length must be 0-dim, [1] or [B]"
            fstride = 1 if (fo.numel() == B and focal.dim() > 0) else 0
        ce = f32c(center)
        bb = f32c(bbox)
        o = _empty(cam, B, H, W, 3)
        d = _empty(cam, B, H, W, 3)
        call("nfi_ray_bundle_fwd", cam, ptr(cam), ptr(fo), fstride, ptr(ce), ptr(bb), B, H, W,
             int(normalize), ptr(o), ptr(d))
        ctx.save_for_backward(cam, fo, ce, bb)
        ctx.meta = (B, H, W, int(normalize), fstride, None if focal is None else focal.shape)
        return o, d

    @staticmethod
    @once_differentiable
    def backward(ctx, go, gd):
        cam, fo, ce, bb = ctx.saved_tensors
        B, H, W, normalize, fstride, fshape = ctx.meta
        go = f32c(go)
        gd = f32c(gd)
        g_cam = _empty(cam, B, 4, 4)
        g_focal = _empty(cam, B)
        call("nfi_ray_bundle_bwd", cam, ptr(cam), ptr(fo), fstride, ptr(ce), ptr(bb), B, H, W, normalize,
             ptr(go), ptr(gd), ptr(g_cam), ptr(g_focal))
        gf = None
        if fo is not None and ctx.needs_input_grad[1]:
            gf = (g_focal if fstride == 1 else g_focal.sum()).reshape(fshape)
        return g_cam, gf, None, None, None, None, None


def ray_bundle(H, W, focal, cam2world, bbox=None, center=None, normalize=False):
    return _RayBundle.apply(cam2world, focal, center, bbox, int(H), int(W), bool(normalize))


# ---------------------------------------------------------------------------------------------
# A2  near / far   lib/nerf_utils.py:227-275 (always under no_grad in the reference, run.py:197)
# ---------------------------------------------------------------------------------------------
_NO_HIT_MSG = ("compute_near_far_planes: no ray intersects the scene box "
               "(min(): expected a non-empty tensor)")


class _Pending(threading.local):
    """Device hit counters of deferred checks (read back in one copy).  Per host thread: the worker
    threads of nn.DataParallel (run.py:636-644) each render on their own device and resolve their
    own checks, so one thread never consumes -- or raises for -- another replica's counters."""

    def __init__(self):
        self.hits = []


_PENDING = _Pending()


def near_far(origins, dirs, scene_range, check_hits=True):
    """check_hits: True -- raise right away like the reference (one host sync, nerf_utils.py:260);
    "defer" -- keep the device counter; `raise_if_no_hits()` reads all pending counters back in one
    copy at the caller's next synchronisation point (pipelined callers: a device->host copy on the
    compute stream would queue behind the bulk output copies on the copy engine); False -- no check."""
    o = f32c(origins.detach())
    d = f32c(dirs.detach())
    shape = o.shape[:-1]
    R = o.numel() // 3
    near = _empty(o, *shape)
    far = _empty(o, *shape)
    ws = _empty(o, NEARFAR_WS_BYTES, dtype=torch.uint8)
    call("nfi_near_far", o, ptr(o), ptr(d), R, float(scene_range), ptr(near), ptr(far), ptr(ws))
    if check_hits == "defer":
        _PENDING.hits.append(ws[:4].view(torch.int32))
    elif check_hits:
        hits = int(ws[:4].view(torch.int32).item())
        if hits == 0:
            # the reference raises from near_plane[mask].min() on an empty tensor (nerf_utils.py:260)
            raise RuntimeError(_NO_HIT_MSG)
    return near, far


def count_launches():
    """Kernels of this library launched so far (graph capture counts once: nfi_b200.graph)."""
    return _lib.LAUNCHES[0]


def take_pending_hits():
    """Hand the deferred hit counters to a caller that resolves them itself (nfi_b200.graph)."""
    pending = list(_PENDING.hits)
    del _PENDING.hits[:]
    return pending


def raise_if_no_hits():
    """Resolve the deferred checks of near_far(check_hits="defer")."""
    if not _PENDING.hits:
        return
    pending = list(_PENDING.hits)
    del _PENDING.hits[:]
    by_dev = {}
    for t in pending:
        by_dev.setdefault(t.device, []).append(t)
    for ts in by_dev.values():
        if int(torch.cat(ts).min().item()) == 0:
            raise RuntimeError(_NO_HIT_MSG)


# ---------------------------------------------------------------------------------------------
# A3  stratified depths + query points   lib/nerf_utils.py:96-122
# ---------------------------------------------------------------------------------------------
def stratified(near, far, S, randomize, u=None, seed=0, offset=0):
    n = f32c(near)
    f = f32c(far)
    R = n.numel()
    z = _empty(n, *n.shape, S)
    uu = f32c(u)
    if uu is not None:
        assert uu.numel() == R * S
    call("nfi_stratified", n, ptr(n), ptr(f), ptr(uu), R, int(S), int(bool(randomize)), int(seed),
         int(offset), ptr(z))
    return z


class _Points(Function):
    @staticmethod
    def forward(ctx, origins, dirs, z):
        o, d, zz = f32c(origins), f32c(dirs), f32c(z)
        S = zz.shape[-1]
        R = zz.numel() // S
        pts = _empty(o, *zz.shape, 3)
        call("nfi_points_fwd", o, ptr(o), ptr(d), ptr(zz), R, S, ptr(pts))
        ctx.save_for_backward(zz)
        ctx.oshape = o.shape
        return pts

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        (zz,) = ctx.saved_tensors
        g = f32c(g)
        S = zz.shape[-1]
        R = zz.numel() // S
        go = _empty(zz, *ctx.oshape)
        gd = _empty(zz, *ctx.oshape)
        call("nfi_points_bwd", zz, ptr(g), ptr(zz), R, S, ptr(go), ptr(gd))
        return go, gd, None


def points_on_rays(origins, dirs, z):
    return _Points.apply(origins, dirs, z.detach())


# ---------------------------------------------------------------------------------------------
# layout boundary: [B, 96, Hp, Wp] NCHW (models/generator.py:475-477) <-> [B, 3, Hp, Wp, 32]
# ---------------------------------------------------------------------------------------------
class _PlanesToCL(Function):
    @staticmethod
    def forward(ctx, planes):
        p = f32c(planes)
        B, Hp, Wp = p.shape[0], p.shape[-2], p.shape[-1]
        assert p.numel() == B * 96 * Hp * Wp, "planes must be [B, 96, Hp, Wp] or [B, 3, 32, Hp, Wp]"
        out = _empty(p, B, 3, Hp, Wp, 32)
        call("nfi_planes_to_cl", p, ptr(p), B, Hp, Wp, ptr(out))
        ctx.shape = planes.shape
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        g = f32c(g)
        B, _, Hp, Wp, _ = g.shape
        out = _empty(g, *ctx.shape)
        call("nfi_planes_grad_from_cl", g, ptr(g), B, Hp, Wp, ptr(out))
        return out


def planes_to_channels_last(planes):
    return _PlanesToCL.apply(planes)


# ---------------------------------------------------------------------------------------------
# A4-A6  triplane gather + decoder MLP + density/colour heads, fused
#        models/generator.py:288-331, 587-681; models/stylegan.py:148-180
# ---------------------------------------------------------------------------------------------
class FieldConfig:
    """Static description of the decoder heads (mirrors NfiField's scalar members)."""

    def __init__(self, attention_values=10, use_sdf=True, scene_range=1.0):
        self.attention_values = int(attention_values)
        self.use_sdf = bool(use_sdf)
        self.scene_range = float(scene_range)


def _param_sizes(n_out):
    return [64 * 32, 64, n_out * 64, n_out, 1, 1]


class PlaneGradSink:
    """One plane-gradient buffer shared by the field passes of one render (coarse + fine read the same
    planes).  Without it each pass zero-fills its own [B,3,Hp,Wp,32] gradient and autograd adds the
    two -- 0.8 GB of fills and a 1.2 GB add per config-2 step.  With it the pass whose backward runs
    first allocates the buffer and returns it as the gradient of `planes_cl`; the others accumulate
    into it with the same atomics (stream-ordered, before the node that consumes it can run: autograd
    keeps the first incoming gradient by reference and waits for every dependent) and return None.
    State is per backward pass (`_current_graph_task_id`), so retained graphs and partial backwards
    start from a fresh buffer."""

    def __init__(self):
        self.buf = None
        self.task = None

    def take(self, like):
        """-> (buffer to accumulate into, gradient to return for planes_cl)"""
        task = torch._C._current_graph_task_id()
        if self.buf is not None and self.task == task and task != -1 and self.buf.shape == like.shape:
            return self.buf, None
        self.buf, self.task = torch.zeros_like(like), task
        return self.buf, self.buf


class _Field(Function):
    """inputs: planes_cl [B,3,Hp,Wp,32], palette [B,A,3]|None, W0,b0,W1,b1,beta,alpha,
    then EITHER (origins, dirs [B*npi,3], z [B*npi,S]) OR points [B,n,3].
    mode 0 -> out[..., 4] = (sigma, r, g, b); mode 1 -> raw decoder output [..., n_out]."""

    @staticmethod
    def forward(ctx, planes_cl, palette, W0, b0, W1, b1, beta, alpha, origins, dirs, z, points,
                cfg, mode, img_w, skip_oob, want_probs, sink=None, planes_bf16=None, probs_grad=False):
        ctx.sink = sink
        ctx.probs_grad = bool(probs_grad)
        pl = f32c(planes_cl)
        # bf16 mode: the kernels gather from `planes_bf16` (a bfloat16 copy of planes_cl, same layout);
        # planes_cl stays the autograd handle, so the plane gradient is fp32 like the master copy
        if planes_bf16 is not None:
            assert planes_bf16.dtype == torch.bfloat16 and planes_bf16.shape == pl.shape \
                and planes_bf16.is_contiguous() and planes_bf16.device == pl.device
        src = pl if planes_bf16 is None else planes_bf16
        B, _, Hp, Wp, C = pl.shape
        assert C == 32 and pl.shape[1] == 3
        pal = f32c(palette)
        field, keep = _lib.make_field_struct(W0, b0, W1, b1, beta, alpha, cfg.attention_values,
                                             cfg.use_sdf, cfg.scene_range, planes_bf16 is not None)
        n_out = field.n_out
        if points is not None:
            pts = f32c(points)
            npi = pts.numel() // 3 // B
            S = 1
            o = d = zz = None
            lead = pts.shape[:-1]
        else:
            o, d, zz = f32c(origins), f32c(dirs), f32c(z)
            S = zz.shape[-1]
            npi = zz.numel() // S // B
            pts = None
            lead = zz.shape
        width = 4 if mode == 0 else n_out
        out = _empty(pl, *lead, width)
        probs = None
        if want_probs and mode == 0 and cfg.attention_values > 0:
            probs = _empty(pl, *lead, cfg.attention_values)
        call("nfi_field_fwd", pl, field, ptr(src), ptr(pal), B, Hp, Wp, ptr(o), ptr(d), ptr(zz), ptr(pts),
             npi, S, mode, int(img_w), int(skip_oob), ptr(out), ptr(probs))
        ctx.save_for_backward(pl, pal, *keep[:4], keep[4], keep[5], o, d, zz, pts)
        ctx.planes_bf16 = planes_bf16
        ctx.cfg, ctx.mode, ctx.img_w, ctx.npi, ctx.S = cfg, mode, int(img_w), npi, S
        ctx.oshape = None if origins is None else origins.shape
        ctx.pshape = None if points is None else points.shape
        ctx.probs = None
        if probs is not None:
            if ctx.probs_grad:
                ctx.probs = probs          # attention probabilities with a gradient (lib/nerf_utils.py:152-155)
            else:
                ctx.mark_non_differentiable(probs)
            return out, probs
        return out, None

    @staticmethod
    @once_differentiable
    def backward(ctx, g_out, g_probs):
        pl, pal, W0, b0, W1, b1, beta, alpha, o, d, zz, pts = ctx.saved_tensors
        cfg, mode = ctx.cfg, ctx.mode
        if g_out is None:
            lead = zz.shape if pts is None else pts.shape[:-1]
            g_out = torch.zeros(*lead, 4 if mode == 0 else W1.shape[0], dtype=torch.float32, device=pl.device)
        need = ctx.needs_input_grad
        want_w = any(need[2:6])           # generator training (run.py:980-986)
        B, _, Hp, Wp, _ = pl.shape
        field, keep = _lib.make_field_struct(W0, b0, W1, b1, beta, alpha, cfg.attention_values,
                                             cfg.use_sdf, cfg.scene_range, ctx.planes_bf16 is not None)
        src = pl if ctx.planes_bf16 is None else ctx.planes_bf16
        n_out = field.n_out
        g = f32c(g_out)
        g_planes = ret_planes = None
        if need[0]:
            if ctx.sink is not None:
                g_planes, ret_planes = ctx.sink.take(pl)
            else:
                g_planes = ret_planes = torch.zeros_like(pl)
        g_pal = torch.zeros_like(pal) if (pal is not None and need[1] and mode == 0) else None
        want_pts = (need[8] or need[9]) if pts is None else need[11]
        g_points = _empty(pl, *g.shape[:-1], 3) if want_pts else None
        want_params = want_w or (cfg.use_sdf and mode == 0 and (need[6] or need[7]))
        g_params = torch.zeros(sum(_param_sizes(n_out)), dtype=torch.float32, device=pl.device) \
            if want_params else None
        if want_w and pts is None and mode == 0:
            _field_bwd_train(field, src, pal, B, Hp, Wp, o, d, zz, ctx.npi, ctx.S, ctx.img_w, g, g_planes, g_pal,
                             g_points, g_params)
        else:
            call("nfi_field_bwd", pl, field, ptr(src), ptr(pal), B, Hp, Wp, ptr(o), ptr(d), ptr(zz), ptr(pts),
                 ctx.npi, ctx.S, mode, ctx.img_w, ptr(g), ptr(g_planes), ptr(g_pal), ptr(g_points),
                 ptr(g_params), int(want_w))
        if g_probs is not None and ctx.probs is not None:
            # gradient arriving at the attention probabilities (semantic map): softmax backward to the logits, then ONE
            # more launch of the field backward in raw-decoder mode with that as the upstream gradient of the decoder
            # output -- same planes / weights / points targets (the point gradient is overwritten by the kernel: summed here)
            pr = ctx.probs
            gpz = f32c(g_probs)
            dlog = pr * (gpz - (pr * gpz).sum(-1, keepdim=True))
            gy2 = torch.cat((torch.zeros_like(dlog[..., :1]), dlog), dim=-1).contiguous()
            g_points2 = torch.empty_like(g_points) if g_points is not None else None
            call("nfi_field_bwd", pl, field, ptr(src), None, B, Hp, Wp, ptr(o), ptr(d), ptr(zz), ptr(pts),
                 ctx.npi, ctx.S, 1, ctx.img_w, ptr(gy2), ptr(g_planes), None, ptr(g_points2),
                 ptr(g_params) if want_w else None, int(want_w))
            if g_points is not None:
                g_points = g_points + g_points2
        go = gd = gp = None
        if want_pts:
            if pts is None:
                R = zz.numel() // ctx.S
                go = _empty(pl, *ctx.oshape)
                gd = _empty(pl, *ctx.oshape)
                call("nfi_points_bwd", pl, ptr(g_points), ptr(zz), R, ctx.S, ptr(go), ptr(gd))
            else:
                gp = g_points.view(ctx.pshape)
        g_beta = g_alpha = gW0 = gb0 = gW1 = gb1 = None
        if g_params is not None:
            sz = _param_sizes(n_out)
            base = sum(sz[:4])
            if cfg.use_sdf and mode == 0:
                g_beta = g_params[base:base + 1].view_as(beta) if need[6] else None
                g_alpha = g_params[base + 1:base + 2].view_as(alpha) if need[7] else None
            if want_w:
                o0, o1, o2 = sz[0], sz[0] + sz[1], sz[0] + sz[1] + sz[2]
                gW0 = g_params[:o0].view_as(W0) if need[2] else None
                gb0 = g_params[o0:o1].view_as(b0) if need[3] else None
                gW1 = g_params[o1:o2].view_as(W1) if need[4] else None
                gb1 = g_params[o2:base].view_as(b1) if need[5] else None
        return (ret_planes, g_pal, gW0, gb0, gW1, gb1, g_beta, g_alpha, go, gd, None, gp,
                None, None, None, None, None, None, None, None)


TRAIN_WORKSPACE_CAP = 6 << 30     # bytes; larger batches go image chunk by image chunk (gradients accumulate)


def _field_bwd_train(field, pl, pal, B, Hp, Wp, o, d, zz, npi, S, img_w, g, g_planes, g_pal, g_points, g_params):
    """Generator-training gradient set (run.py:980-986) through nfi_field_bwd_train: the warp-specialised backward
    plus one weight-gradient reduction kernel.  The workspace (705 bytes per sample) is bounded by splitting
    the batch by image; every output slice of an image chunk is contiguous."""
    l = _lib.lib()
    per_img = int(l.nfi_field_bwd_train_workspace_bytes(npi * S))
    nb = max(1, min(B, TRAIN_WORKSPACE_CAP // max(per_img, 1)))
    ws_bytes = int(l.nfi_field_bwd_train_workspace_bytes(nb * npi * S))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=pl.device)
    o2, d2 = o.reshape(B, npi, 3), d.reshape(B, npi, 3)
    z2, g2 = zz.reshape(B, npi, S), g.reshape(B, npi, S, 4)
    gp2 = None if g_points is None else g_points.view(B, npi, S, 3)
    for b0 in range(0, B, nb):
        b1 = min(B, b0 + nb)
        sl = slice(b0, b1)
        call("nfi_field_bwd_train", pl, field, ptr(pl[sl]), None if pal is None else ptr(pal[sl]), b1 - b0, Hp, Wp,
             ptr(o2[sl]), ptr(d2[sl]), ptr(z2[sl]), npi, S, img_w, ptr(g2[sl]),
             None if g_planes is None else ptr(g_planes[sl]), None if g_pal is None else ptr(g_pal[sl]),
             None if gp2 is None else ptr(gp2[sl]), ptr(g_params), ptr(ws), ws_bytes)


# ---------------------------------------------------------------------------------------------
# N2  regulariser path: decoder output with a DIFFERENTIABLE coordinate gradient (double backward)
#     models/generator.py:505-585, lib/ops.py:58-120 (grid_sample2d), TriplanarDecoder.forward(requires_double_backward=True)
# ---------------------------------------------------------------------------------------------
class _DecoderGradX(Function):
    """gx = J^T gy: the gradient of the raw decoder output w.r.t. the query coordinates (box units), as the
    reference obtains it with autograd.grad(..., create_graph=True) (generator.py:533-536).  Its backward -- the
    double backward -- gives the gradients of any function of gx w.r.t. planes, W0, b0, W1 (hand-written kernel,
    csrc/nfi_sdfgrad.cu).  No gradient flows to the coordinates or to gy from here (the reference's losses need none)."""

    @staticmethod
    def forward(ctx, planes_cl, W0, b0, W1, coords, gy):
        pl, x, g = f32c(planes_cl), f32c(coords), f32c(gy)
        B, _, Hp, Wp, _ = pl.shape
        n = x.numel() // 3 // B
        n_out = W1.shape[0]
        field, keep = _lib.make_field_struct(W0, b0, W1, torch.zeros(n_out, device=pl.device), None, None,
                                             n_out - 1 if n_out != 4 else 0, False, 1.0)
        gx = _empty(pl, *x.shape)
        call("nfi_decoder_gradx_fwd", pl, field, ptr(pl), B, Hp, Wp, ptr(x), ptr(g), n, ptr(gx))
        ctx.save_for_backward(pl, keep[0], keep[1], keep[2], x, g)
        return gx

    @staticmethod
    @once_differentiable
    def backward(ctx, gg):
        pl, W0, b0, W1, x, g = ctx.saved_tensors
        B, _, Hp, Wp, _ = pl.shape
        n = x.numel() // 3 // B
        n_out = W1.shape[0]
        need = ctx.needs_input_grad
        field, keep = _lib.make_field_struct(W0, b0, W1, torch.zeros(n_out, device=pl.device), None, None,
                                             n_out - 1 if n_out != 4 else 0, False, 1.0)
        gpl = torch.zeros_like(pl) if need[0] else None
        gW0 = torch.zeros_like(W0) if need[1] else None
        gb0 = torch.zeros_like(b0) if need[2] else None
        gW1 = torch.zeros_like(W1) if need[3] else None
        call("nfi_decoder_gradx_bwd", pl, field, ptr(pl), B, Hp, Wp, ptr(x), ptr(g), ptr(f32c(gg)), n, ptr(gpl),
             ptr(gW0), ptr(gb0), ptr(gW1))
        return gpl, gW0, gb0, gW1, None, None


class _DecoderDD(Function):
    """Raw decoder output at explicit points whose backward is itself differentiable in the one way the reference
    uses: the gradient w.r.t. the coordinates (`_DecoderGradX`).  The gradients w.r.t. planes and weights are the
    ordinary first-order ones (nfi_field_bwd, raw-decoder mode)."""

    @staticmethod
    def forward(ctx, planes_cl, W0, b0, W1, b1, coords):
        n_out = W1.shape[0]
        cfg = FieldConfig(attention_values=(n_out - 1) if n_out != 4 else 0, use_sdf=False, scene_range=1.0)
        with torch.no_grad():
            y, _ = field_on_points(planes_cl, None, W0, b0, W1, b1, None, None, coords, cfg, mode=1)
        ctx.save_for_backward(planes_cl, W0, b0, W1, b1, coords)
        ctx.cfg = cfg
        return y

    @staticmethod
    def backward(ctx, gy):
        planes_cl, W0, b0, W1, b1, coords = ctx.saved_tensors
        need = ctx.needs_input_grad
        g_coords = None
        if need[5]:
            g_coords = _DecoderGradX.apply(planes_cl, W0, b0, W1, coords, gy)
        grads = [None] * 5
        if any(need[:5]):
            with torch.enable_grad():
                leaves = [t.detach().requires_grad_(need[i]) for i, t in enumerate((planes_cl, W0, b0, W1, b1))]
                y, _ = field_on_points(leaves[0], None, leaves[1], leaves[2], leaves[3], leaves[4], None, None,
                                       coords.detach(), ctx.cfg, mode=1)
                wanted = [t for t, nd in zip(leaves, need[:5]) if nd]
                got = torch.autograd.grad(y, wanted, gy.detach())
            it = iter(got)
            grads = [next(it) if nd else None for nd in need[:5]]
        return (*grads, g_coords)


def decoder_double_backward(planes_cl, W0, b0, W1, b1, coords):
    """y [B,N,n_out] = decoder(coords) with a backward that supports create_graph=True w.r.t. `coords`."""
    return _DecoderDD.apply(planes_cl, W0, b0, W1, b1, coords)


def decoder_mlp(feat, W0, b0, W1, b1, want_hidden=False):
    """Decoder MLP alone on the tensor cores (forward only): feat [n,32] -> y [n,n_out]."""
    ft = f32c(feat.detach())
    n = ft.shape[0]
    n_out = W1.shape[0]
    field, keep = _lib.make_field_struct(W0, b0, W1, b1, None, None, n_out - 1 if n_out != 4 else 0, False, 1.0)
    y = _empty(ft, n, n_out)
    hp = _empty(ft, n, 64) if want_hidden else None
    call("nfi_decoder_mlp_fwd", ft, field, ptr(ft), n, ptr(y), ptr(hp))
    return (y, hp) if want_hidden else y


def planes_bf16_copy(planes_cl):
    """bf16 mode: the bfloat16 copy of channels-last planes the kernels gather from (round to nearest
    even, like Tensor.bfloat16()); pass it as `planes_bf16=` next to the fp32 planes_cl."""
    return planes_cl.detach().to(torch.bfloat16).contiguous()


def field_on_rays(planes_cl, palette, W0, b0, W1, b1, beta, alpha, origins, dirs, z, cfg,
                  img_w=0, skip_oob=False, want_probs=False, sink=None, planes_bf16=None, probs_grad=False):
    """(sigma, r, g, b) [.., S, 4] for the samples o + d*z of every ray (no points tensor in HBM).
    `sink`: a PlaneGradSink shared with the other passes over the same planes_cl."""
    return _Field.apply(planes_cl, palette, W0, b0, W1, b1, beta, alpha, origins, dirs, z.detach(),
                        None, cfg, 0, img_w, skip_oob, want_probs, sink, planes_bf16, probs_grad)


def field_on_points(planes_cl, palette, W0, b0, W1, b1, beta, alpha, points, cfg, mode=0,
                    want_probs=False, planes_bf16=None):
    """Heads (mode 0) or raw decoder output (mode 1) at explicit points [B, n, 3]."""
    return _Field.apply(planes_cl, palette, W0, b0, W1, b1, beta, alpha, None, None, None, points,
                        cfg, mode, 0, False, want_probs, None, planes_bf16)


def _sdf_normals(planes_cl, W0, b0, W1, b1, cfg, origins, dirs, z, points, img_w):
    with torch.no_grad():
        assert cfg.use_sdf, "normals need the SDF head (models/generator.py:600)"
        pl = f32c(planes_cl)
        B, _, Hp, Wp, _ = pl.shape
        if points is None:
            o, d, zz = f32c(origins.detach()), f32c(dirs.detach()), f32c(z.detach())
            pts, S, lead = None, zz.shape[-1], zz.shape
            npi = zz.numel() // S // B
        else:
            pts = f32c(points.detach())
            o = d = zz = None
            S, lead, npi = 1, pts.shape[:-1], pts.numel() // 3 // B
        # raw-decoder mode: no heads, so the struct carries no beta / alpha
        field, keep = _lib.make_field_struct(W0.detach(), b0.detach(), W1.detach(), b1.detach(), None, None,
                                             cfg.attention_values, False, cfg.scene_range)
        g = torch.zeros(*lead, field.n_out, dtype=torch.float32, device=pl.device)
        g[..., 0] = 1.0
        g_points = _empty(pl, *lead, 3)
        call("nfi_field_bwd", pl, field, ptr(pl), None, B, Hp, Wp, ptr(o), ptr(d), ptr(zz), ptr(pts),
             npi, S, 1, int(img_w), ptr(g), None, None, ptr(g_points), None, 0)
        return torch.nn.functional.normalize(g_points, dim=-1)


def sdf_normals_on_rays(planes_cl, W0, b0, W1, b1, origins, dirs, z, cfg, img_w=0):
    """Unit SDF gradients at the samples o + d*z of every ray, [.., S, 3] -- the sampler's 'normals'
    output (models/generator.py:599-602, 614-618: autograd of the distance channel wrt the query
    points, then F.normalize).  Forward-only, as in the reference (it detaches everything in this
    mode).  One launch of the fused field backward in raw-decoder mode with the upstream gradient
    e_0 (distance channel) and no plane-gradient target: d y_0 / d point comes out of the same
    tcgen05 chain (dY -> dH -> dF) and the bilinear tangents the pose gradient uses."""
    return _sdf_normals(planes_cl, W0, b0, W1, b1, cfg, origins, dirs, z, None, img_w)


def sdf_normals_on_points(planes_cl, W0, b0, W1, b1, points, cfg):
    """Same at explicit world-space points [B, n, 3] -> [B, n, 3]."""
    return _sdf_normals(planes_cl, W0, b0, W1, b1, cfg, None, None, None, points, 0)


# ---------------------------------------------------------------------------------------------
# A7 / A8  coarse weights, EG3D smoothing, inverse-CDF sampling (all under no_grad in the
#          reference: run.py:261-281)
# ---------------------------------------------------------------------------------------------
def weights_only(sigma, dirs, z, smooth=False, sigma_stride=1):
    """sigma [.., S] (or a strided view base with `sigma_stride` floats between samples)."""
    zz = f32c(z.detach())
    S = zz.shape[-1]
    R = zz.numel() // S
    sg = f32c(sigma.detach())
    d = f32c(dirs.detach())
    w = _empty(zz, *zz.shape)
    call("nfi_weights_only", zz, ptr(sg), int(sigma_stride), ptr(d), ptr(zz), R, S, int(smooth), ptr(w))
    return w


def sample_pdf(bins, weights, n, u=None, return_indices=False):
    b = f32c(bins.detach())
    w = f32c(weights.detach())
    R, M = b.shape
    assert w.shape == (R, M - 1)
    uu = f32c(u)
    out = _empty(b, R, n)
    inds = _empty(b, R, n, dtype=torch.int32) if return_indices else None
    call("nfi_sample_pdf", b, ptr(b), ptr(w), R, M, int(n), ptr(uu), ptr(out), ptr(inds))
    return (out, inds) if return_indices else out


def importance(sigma, dirs, z_coarse, u=None, sigma_stride=1):
    """run.py:261-281 in one kernel: coarse sigma + depths -> S fine depths per ray."""
    zc = f32c(z_coarse.detach())
    S = zc.shape[-1]
    R = zc.numel() // S
    sg = f32c(sigma.detach())
    d = f32c(dirs.detach())
    uu = f32c(u)
    zf = _empty(zc, *zc.shape)
    call("nfi_importance", zc, ptr(sg), int(sigma_stride), ptr(d), ptr(zc), R, S, ptr(uu), ptr(zf))
    return zf


def cumprod_exclusive(x):
    """lib/nerf_utils.py:20-25 (forward only)."""
    xx = f32c(x.detach())
    n = xx.shape[-1]
    out = torch.empty_like(xx)
    call("nfi_cumprod_exclusive", xx, ptr(xx), xx.numel() // n, n, ptr(out))
    return out


# ---------------------------------------------------------------------------------------------
# A10  alpha compositing   lib/nerf_utils.py:125-163
# ---------------------------------------------------------------------------------------------
class _Composite(Function):
    @staticmethod
    def forward(ctx, sigma, rgb, dirs, z, normals, semantics, white):
        sg, col, d, zz = f32c(sigma), f32c(rgb), f32c(dirs), f32c(z)
        nrm, sem = f32c(normals), f32c(semantics)
        S = zz.shape[-1]
        R = zz.numel() // S
        lead = zz.shape[:-1]
        n_sem = 0 if sem is None else sem.shape[-1]
        rgb_map = _empty(zz, *lead, 3)
        depth = _empty(zz, *lead)
        mask = _empty(zz, *lead)
        nmap = _empty(zz, *lead, 3) if nrm is not None else None
        smap = _empty(zz, *lead, n_sem) if sem is not None else None
        call("nfi_composite_fwd", zz, ptr(sg), ptr(col), ptr(d), ptr(zz), ptr(nrm), ptr(sem), n_sem, R, S,
             int(white), ptr(rgb_map), ptr(depth), ptr(mask), ptr(nmap), ptr(smap))
        ctx.save_for_backward(sg, col, d, zz, sem)
        ctx.white = int(white)
        ctx.mark_non_differentiable(depth)
        if nmap is not None:
            ctx.mark_non_differentiable(nmap)   # gradient to `normals` is not produced (N1, next)
        return rgb_map, depth, mask, nmap, smap

    @staticmethod
    @once_differentiable
    def backward(ctx, g_rgb_map, _g_depth, g_mask, _g_nmap, g_smap):
        sg, col, d, zz, sem = ctx.saved_tensors
        S = zz.shape[-1]
        R = zz.numel() // S
        n_sem = 0 if sem is None else sem.shape[-1]
        g_rgb_map = f32c(g_rgb_map)
        g_mask = f32c(g_mask)
        g_smap = f32c(g_smap) if sem is not None else None
        if g_rgb_map is None:
            g_rgb_map = torch.zeros(*zz.shape[:-1], 3, dtype=torch.float32, device=zz.device)
        g_sigma = _empty(zz, *zz.shape)
        g_rgb = _empty(zz, *col.shape) if ctx.needs_input_grad[1] else None
        g_dirs = _empty(zz, *d.shape) if ctx.needs_input_grad[2] else None
        g_sem = torch.zeros_like(sem) if (sem is not None and ctx.needs_input_grad[5]) else None
        call("nfi_composite_bwd", zz, ptr(sg), ptr(col), ptr(d), ptr(zz), ptr(sem), n_sem, R, S, ctx.white,
             ptr(g_rgb_map), ptr(g_mask), ptr(g_smap), ptr(g_sigma), ptr(g_rgb), ptr(g_dirs), ptr(g_sem))
        return g_sigma, g_rgb, g_dirs, None, None, g_sem, None


def composite(sigma, rgb, dirs, z, normals=None, semantics=None, white_background=True):
    return _Composite.apply(sigma, rgb, dirs, z.detach(), normals, semantics, bool(white_background))


# ---------------------------------------------------------------------------------------------
# A9 + A10 fused: depth merge of coarse + fine samples and compositing   run.py:283-348
# ---------------------------------------------------------------------------------------------
class _MergeComposite(Function):
    @staticmethod
    def forward(ctx, coarse, fine, z_coarse, z_fine, dirs, white, cam2world):
        c, f = f32c(coarse), f32c(fine)
        zc, zf, d = f32c(z_coarse), f32c(z_fine), f32c(dirs)
        S = zc.shape[-1]
        R = zc.numel() // S
        lead = zc.shape[:-1]
        cam = f32c(cam2world)
        rpi = R // cam.shape[0] if cam is not None else 0
        rgb_map = _empty(zc, *lead, 3)
        depth = _empty(zc, *lead)
        mask = _empty(zc, *lead)
        perm = _empty(zc, *lead, 2 * S, dtype=torch.uint8)
        zs = _empty(zc, *lead, 2 * S)
        call("nfi_merge_composite_fwd", zc, ptr(c), ptr(f), ptr(zc), ptr(zf), ptr(d), R, S, int(white),
             ptr(cam), rpi, ptr(rgb_map), ptr(depth), ptr(mask), ptr(perm), ptr(zs))
        ctx.save_for_backward(c, f, zs, perm, d)
        ctx.white = int(white)
        ctx.mark_non_differentiable(depth, perm, zs)
        return rgb_map, depth, mask, perm, zs

    @staticmethod
    @once_differentiable
    def backward(ctx, g_rgb_map, _gd, g_mask, _gp, _gz):
        c, f, zs, perm, d = ctx.saved_tensors
        S = c.shape[-2]
        R = c.numel() // (4 * S)
        g_rgb_map = f32c(g_rgb_map)
        g_mask = f32c(g_mask)
        g_c = torch.empty_like(c)
        g_f = torch.empty_like(f)
        call("nfi_merge_composite_bwd", c, ptr(c), ptr(f), ptr(zs), ptr(perm), ptr(d), R, S, ctx.white,
             ptr(g_rgb_map), ptr(g_mask), ptr(g_c), ptr(g_f))
        # d|dir| is not returned: render() passes unit directions, whose norm has a zero
        # gradient after F.normalize's projection (run.py:196).
        return g_c, g_f, None, None, None, None, None


def merge_composite(coarse, fine, z_coarse, z_fine, dirs, white_background, cam2world_for_zbuffer=None):
    """coarse / fine: (sigma, r, g, b) [.., S, 4].  Returns rgb, depth, mask, perm (uint8), z_sorted."""
    return _MergeComposite.apply(coarse, fine, z_coarse.detach(), z_fine.detach(), dirs.detach(),
                                 bool(white_background), cam2world_for_zbuffer)
