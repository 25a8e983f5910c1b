"""The fused render: `run.py::render` (run.py:176-350; the six script-local copies, the z-buffer
variant at eval_kitti_persp.py:221-228) rebuilt ray-major on the sm_100a kernels.

Two entry points:
  * `render_planes(...)`  -- the hot path proper, on explicit tensors (planes, attention values,
    decoder weights, camera).  Query points, gathered features, hidden activations and per-sample
    sort indices never reach HBM; what does: rays, depths, one float4 per sample, the images.
  * `make_render(args, dataset_config)` -> `render(target_model, height, width, ...)` with the
    reference's exact signature and 6-tuple return, for `ParallelModel.forward` (run.py:597-611).
"""
from __future__ import annotations

import contextlib
from typing import Optional

import torch

from . import ops
from .decoder import Field, _as_planes_tensor


def render_planes(planes, palette, field: Field, height: int, width: int, tform_cam2world,
                  focal_length=None, center=None, bbox=None, depth_samples_per_ray: int = 64,
                  randomize: bool = True, white_background: bool = False, fine_sampling: bool = True,
                  force_no_cam_grad: bool = False, zbuffer_depth: bool = False,
                  planes_layout: str = "nchw", u_strat=None, u_pdf=None, check_hits=True,
                  return_aux: bool = False, compute_normals: bool = False,
                  compute_semantics: bool = False, compute_coords: bool = False,
                  plane_dtype: torch.dtype = torch.float32, z_fine=None):
    """Render `B` images.  planes: [B,96,Hp,Wp] or [B,3,32,Hp,Wp] (layout 'nchw', the StyleGAN2
    output of models/generator.py:475-477) or [B,3,Hp,Wp,32] ('cl').  palette: attention values
    [B,A,3] (None for the plain RGB head).  Returns (rgb [B,H,W,3], depth [B,H,W], mask [B,H,W])
    (+ aux dict).  `u_strat` [B,H,W,S] / `u_pdf` [B*H*W,S] override the uniforms that are
    otherwise drawn with torch.rand in the reference's order (nerf_utils.py:115, :204).
    compute_normals / compute_semantics / compute_coords (run.py:186-188) add the
    maps aux["normals"] [B,H,W,3] (forward-only, as the reference detaches them) and aux["semantics"] [B,H,W,A]
    (coords [B,H,W,3] take the semantics slot, run.py:337-338), the latter with gradients when autograd is on
    (lib/nerf_utils.py:152-155); they imply `return_aux`.
    plane_dtype=torch.bfloat16 is the bf16 mode of the contract (1e-2 instead of 1e-4): the gather reads
    a bfloat16 copy of the planes (half the texel bytes); everything after the load, and every gradient,
    stays fp32.  Not available together with the extra maps.
    z_fine [B,H,W,S] (test hook): use these fine depths instead of running A7 + A8 -- lets a parity test
    feed the reference's own importance samples, so that what remains is the differentiable path alone."""
    H, W, S = int(height), int(width), int(depth_samples_per_ray)
    cam = tform_cam2world
    B = cam.shape[0]
    dev = cam.device
    # A1 (+ F.normalize, run.py:196), A2 (no_grad, run.py:197-200), A3 depths
    o, d = ops.ray_bundle(H, W, focal_length, cam, bbox, center, normalize=True)
    near, far = ops.near_far(o, d, field.cfg.scene_range, check_hits=check_hits)
    if randomize and u_strat is None:
        u_strat = torch.rand((B, H, W, S), dtype=torch.float32, device=dev)
    z_c = ops.stratified(near, far, S, randomize, u_strat if randomize else None)
    if force_no_cam_grad:                       # run.py:211-214
        o_s, d_s = o.detach(), d.detach()
    else:
        o_s, d_s = o, d
    planes_cl = planes if planes_layout == "cl" else ops.planes_to_channels_last(planes)
    wts = field.tensors()
    assert plane_dtype in (torch.float32, torch.bfloat16), "plane_dtype: torch.float32 or torch.bfloat16"
    pl16 = ops.planes_bf16_copy(planes_cl) if plane_dtype == torch.bfloat16 else None
    if pl16 is not None:
        assert not (compute_normals or compute_semantics or compute_coords), \
            "bf16 mode covers rgb / depth / mask and their gradients"
    # A4-A6 on the coarse samples: (sigma, r, g, b) per sample
    extras = compute_normals or compute_semantics or compute_coords
    want_sem = compute_semantics and not compute_coords
    if compute_semantics:
        assert field.cfg.attention_values > 0      # run.py:229
    sink = ops.PlaneGradSink() if fine_sampling else None     # coarse + fine: one gradient buffer
    # The semantic / coordinate maps carry gradients like the reference's (lib/nerf_utils.py:152-155 keeps the weights
    # and the per-sample payload attached) when autograd is on; the evaluation callers run under no_grad and take the
    # cheaper forward-only route.  Normals stay forward-only (the reference detaches them, generator.py:619-622).
    grad_maps = torch.is_grad_enabled() and (compute_semantics or compute_coords)
    coarse, probs_c = ops.field_on_rays(planes_cl, palette, *wts, o_s, d_s, z_c, field.cfg, img_w=W,
                                        skip_oob=not want_sem, want_probs=want_sem, sink=sink, planes_bf16=pl16,
                                        probs_grad=grad_maps and want_sem)
    nrm_c = ops.sdf_normals_on_rays(planes_cl, *wts[:4], o_s, d_s, z_c, field.cfg, img_w=W) \
        if compute_normals else None
    aux = {"origins": o, "dirs": d, "near": near, "far": far, "z_coarse": z_c, "normals": None,
           "semantics": None}
    if fine_sampling:
        # A7 + A8 (no_grad, run.py:261-281)
        if randomize and u_pdf is None:
            u_pdf = torch.rand((B * H * W, S), dtype=torch.float32, device=dev)
        if z_fine is None:
            z_f = ops.importance(coarse, d, z_c, u_pdf if randomize else None, sigma_stride=4)
        else:
            z_f = z_fine.detach().reshape(z_c.shape).contiguous().float()
        fine, probs_f = ops.field_on_rays(planes_cl, palette, *wts, o_s, d_s, z_f, field.cfg, img_w=W,
                                          skip_oob=not want_sem, want_probs=want_sem, sink=sink, planes_bf16=pl16,
                                          probs_grad=grad_maps and want_sem)
        # A9 + A10
        rgb, depth, mask, perm, z_sorted = ops.merge_composite(
            coarse, fine, z_c, z_f, d, white_background, cam.detach() if zbuffer_depth else None)
        aux.update(z_fine=z_f, perm=perm, z_sorted=z_sorted)
        if extras:
            # N1: per-sample payloads in depth order (run.py:312-335) and their maps
            # (nerf_utils.py:148-161); forward-only, from the stage-level compositing kernel
            with (contextlib.nullcontext() if grad_maps else torch.no_grad()):
                det = (lambda t: t) if grad_maps else (lambda t: t.detach())

                def sort(c, f):
                    both = torch.cat((c, f), dim=-2)
                    return both.gather(-2, perm.long().unsqueeze(-1).expand(*perm.shape, both.shape[-1]))
                nrm = sem = None
                if compute_normals:
                    with torch.no_grad():
                        nrm_f = ops.sdf_normals_on_rays(planes_cl, *wts[:4], o_s, d_s, z_f, field.cfg, img_w=W)
                        nrm = sort(nrm_c, nrm_f)
                if compute_coords:
                    sem = sort(ops.points_on_rays(det(o_s), det(d_s), z_c), ops.points_on_rays(det(o_s), det(d_s), z_f))
                elif compute_semantics:
                    sem = sort(probs_c, probs_f)
                both = sort(det(coarse), det(fine))
                _, _, _, nmap, smap = ops.composite(both[..., 0], both[..., 1:], d.detach(), z_sorted, nrm,
                                                    sem, white_background)
                aux.update(normals=nmap, semantics=smap)
    else:
        rgb, depth, mask, _, _ = ops.composite(coarse[..., 0], coarse[..., 1:], d.detach(), z_c,
                                               None, None, white_background)
        if zbuffer_depth:
            raise NotImplementedError("z-buffer depth without fine sampling")
        if extras:
            with torch.no_grad():
                sem = ops.points_on_rays(o_s.detach(), d_s.detach(), z_c) if compute_coords else probs_c
                _, _, _, nmap, smap = ops.composite(coarse[..., 0].detach(), coarse[..., 1:].detach(),
                                                    d.detach(), z_c, nrm_c, sem, white_background)
                aux.update(normals=nmap, semantics=smap)
    if return_aux or extras:
        return rgb, depth, mask, aux
    return rgb, depth, mask


def render_viewdir(model_for, height, width, tform_cam2world, focal_length=None,
                   center=None, bbox=None, depth_samples_per_ray: int = 64, randomize: bool = True,
                   white_background: bool = False, fine_sampling: bool = True, force_no_cam_grad: bool = False,
                   zbuffer_depth: bool = False, check_hits=True, scene_range: float = 1.0):
    """the view-direction branch (CARLA; models/generator.py:189-253, 468-469, 662-663).

    With `use_viewdir` the decoder has 1 + 32 outputs and the 32 features of a sample are modulated by a per-ray
    vector before the colour head: y = LeakyReLU(x_ray + features) -> EqualizedLinear(32 -> A or 3).  This path is a
    HYBRID, not the fused kernels: rays, depths, importance sampling, merge and compositing are the library's
    kernels; the triplane gather + decoder MLP run on the tensor-core kernel in raw-decoder mode, three passes of 11
    of the 33 outputs each (the kernels hold at most 16 outputs); the per-ray modulation, the output layer and the
    density / colour heads (generator.py:629-636, 661-679) are PyTorch ops on [B, N, 33] tensors.  Functionally
    complete with autograd to planes, decoder, mapper, palette and pose; about 4x the fused path's time.
    `model_for(viewdirs)` is called ONCE, after the rays exist (the generator needs the view directions before it
    can build its sampler, generator.py:468-469), with `viewdirs` [B,H,W,1,3]; it returns (gen, planes [B,96,Hp,Wp],
    palette [B,A,3] or None, mapper_closure) -- `gen` supplies decoder / beta / alpha / scene_range /
    attention_values / use_sdf, `mapper_closure` is the reference's (generator.py:241-251).  `scene_range` must be
    given because the near/far planes are needed before the generator runs."""
    import torch.nn.functional as F
    H, W, S = int(height), int(width), int(depth_samples_per_ray)
    cam = tform_cam2world
    B = cam.shape[0]
    dev = cam.device
    sr = float(scene_range)
    o, d = ops.ray_bundle(H, W, focal_length, cam, bbox, center, normalize=True)
    near, far = ops.near_far(o, d, sr, check_hits=check_hits)
    u = torch.rand((B, H, W, S), dtype=torch.float32, device=dev) if randomize else None
    z_c = ops.stratified(near, far, S, randomize, u)
    o_s, d_s = (o.detach(), d.detach()) if force_no_cam_grad else (o, d)
    gen, planes, palette, closure = model_for(d_s.unsqueeze(-2))
    assert abs(float(gen.scene_range) - sr) < 1e-12
    planes_cl = ops.planes_to_channels_last(planes)
    l0, l2 = gen.decoder.net[0], gen.decoder.net[2]
    n_out = l2.weight.shape[0]
    assert n_out == 33, "use_viewdir decoders have 1 + 32 outputs (models/generator.py:375-376)"
    cfg11 = ops.FieldConfig(attention_values=10, use_sdf=False, scene_range=1.0)
    A = int(gen.attention_values)

    def field(z):
        pts = ops.points_on_rays(o_s, d_s, z)                                   # [B,H,W,S,3]
        x = (pts / sr).reshape(B, -1, 3)
        with torch.no_grad():
            oob = (x.abs() > 1).any(dim=-1).float()                             # generator.py:605-607
        ys = [ops.field_on_points(planes_cl, None, l0.weight, l0.bias, l2.weight[k:k + 11], l2.bias[k:k + 11], None,
                                  None, x, cfg11, mode=1)[0] for k in (0, 11, 22)]
        y = torch.cat(ys, dim=-1)                                               # [B,N,33]
        dist, feat = y[..., 0], y[..., 1:].contiguous()
        if gen.use_sdf:                                                         # generator.py:629-636
            nd = -dist
            cdf = 0.5 + 0.5 * torch.sign(nd) * (1 - torch.exp(-nd.abs() / gen.beta))
            sigma = (1 / gen.alpha) * (cdf * (1 - oob))
        else:                                                                   # generator.py:637-641
            sigma = F.softplus(dist - 1) * (1 - oob)
        feat = closure(feat)                                                    # generator.py:662-663
        if A == 0:
            rgb = torch.sigmoid(feat) * 2.004 - 1.002                           # generator.py:36-39, 665-666
        else:
            rgb = torch.matmul(F.softmax(feat, dim=-1), palette)                # generator.py:668-678
        return torch.cat((sigma.unsqueeze(-1), rgb), dim=-1).view(B, H, W, S, 4).contiguous()

    coarse = field(z_c)
    if not fine_sampling:
        rgb, depth, mask, _, _ = ops.composite(coarse[..., 0], coarse[..., 1:], d.detach(), z_c, None, None,
                                               white_background)
        return rgb, depth, mask
    u_pdf = torch.rand((B * H * W, S), dtype=torch.float32, device=dev) if randomize else None
    z_f = ops.importance(coarse.detach(), d, z_c, u_pdf, sigma_stride=4)
    fine = field(z_f)
    rgb, depth, mask, _, _ = ops.merge_composite(coarse, fine, z_c, z_f, d, white_background,
                                                 cam.detach() if zbuffer_depth else None)
    return rgb, depth, mask


def _closure_vars(fn):
    """Free variables of the reference's `sampler` closure (models/generator.py:587).  Cells that the
    generator never assigned stay out of the result: `viewdir_mapper_closure` only exists with
    use_viewdir (generator.py:468-469), `attention_values` only with attention_values > 0 (:451-463)."""
    out = {}
    for name, cell in zip(fn.__code__.co_freevars, fn.__closure__ or ()):
        try:
            out[name] = cell.cell_contents
        except ValueError:          # empty cell
            pass
    return out


def make_render(args, dataset_config, zbuffer_depth: bool = False):
    """Build a `render` with the reference's signature (run.py:176-191), bound to the script's
    `args` (use_viewdir, use_sdf, attention_values, fine_sampling) and `dataset_config`
    (scene_range, white_background) globals that the reference function reads (run.py:200-348)."""

    def _render_with_viewdir(target_model, height, width, cam, focal, center, bbox, model_input, S, randomize,
                             extra_model_outputs, extra_model_inputs, force_no_cam_grad):
        box = {}

        def model_for(viewdirs):
            outs = target_model(viewdirs, model_input, ["sampler"] + list(extra_model_outputs), extra_model_inputs)
            cv = _closure_vars(outs.pop("sampler"))
            box["outs"] = outs
            gen = cv["self"]
            palette = cv.get("attention_values") if gen.attention_values > 0 else None
            return gen, _as_planes_tensor(cv["xy"], cv["xz"], cv["yz"]), palette, cv["viewdir_mapper_closure"]

        rgb, depth, mask = render_viewdir(model_for, height, width, cam, focal, center, bbox, S, randomize=randomize,
                                          white_background=dataset_config["white_background"],
                                          fine_sampling=args.fine_sampling, force_no_cam_grad=force_no_cam_grad,
                                          zbuffer_depth=zbuffer_depth, scene_range=dataset_config["scene_range"])
        return rgb, depth, mask, None, None, box["outs"]

    def render(target_model, height, width, tform_cam2world, focal_length, center, bbox, model_input,
               depth_samples_per_ray, randomize=True, compute_normals=False, compute_semantics=False,
               compute_coords=False, extra_model_outputs=[], extra_model_inputs={},
               force_no_cam_grad=False):
        if args.use_viewdir:
            if compute_normals or compute_semantics or compute_coords:
                raise NotImplementedError("nfi_b200.render: normals / semantics / coords maps with use_viewdir")
            return _render_with_viewdir(target_model, height, width, tform_cam2world, focal_length, center, bbox,
                                        model_input, depth_samples_per_ray, randomize, extra_model_outputs,
                                        extra_model_inputs, force_no_cam_grad)
        if compute_normals:
            assert args.use_sdf                    # run.py:226
        # The generator still produces planes and attention values (StyleGAN2 synthesis, out of
        # scope); the closure it returns holds them -- read them instead of calling it.
        model_outputs = target_model(None, model_input, ["sampler"] + list(extra_model_outputs),
                                     extra_model_inputs)
        sampler = model_outputs.pop("sampler")
        cv = _closure_vars(sampler)
        gen = cv["self"]
        planes = _as_planes_tensor(cv["xy"], cv["xz"], cv["yz"])
        palette = cv.get("attention_values") if gen.attention_values > 0 else None
        field = Field.from_generator(gen)
        assert abs(float(gen.scene_range) - float(dataset_config["scene_range"])) < 1e-12
        rgb, depth, mask, aux = render_planes(
            planes, palette, field, height, width, tform_cam2world, focal_length, center, bbox,
            depth_samples_per_ray, randomize=randomize,
            white_background=dataset_config["white_background"], fine_sampling=args.fine_sampling,
            force_no_cam_grad=force_no_cam_grad, zbuffer_depth=zbuffer_depth, return_aux=True,
            compute_normals=compute_normals, compute_semantics=compute_semantics,
            compute_coords=compute_coords)
        return rgb, depth, mask, aux["normals"], aux["semantics"], model_outputs

    return render
