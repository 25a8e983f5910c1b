"""Drop-in for `TriplanarDecoder.forward` and the `sampler` closure of the reference
(models/generator.py:288-331, 587-681).  The decoder owns no parameters: weights are read from the
reference `nn.Module` at call time, so checkpoints (`decoder.net.0.weight` ... `beta`, `alpha`)
keep loading unchanged and gradients flow to the module's own parameters.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .ops import FieldConfig


class Field:
    """Decoder weights + head configuration, gathered from a reference Generator or given directly."""

    def __init__(self, W0, b0, W1, b1, beta=None, alpha=None, scene_range=1.0, attention_values=10,
                 use_sdf=True):
        self.W0, self.b0, self.W1, self.b1 = W0, b0, W1, b1
        self.beta, self.alpha = beta, alpha
        self.cfg = FieldConfig(attention_values, use_sdf, scene_range)
        n_out = W1.shape[0]
        want = 1 + (attention_values if attention_values > 0 else 3)
        assert n_out == want, "decoder.net[2] has %d outputs, heads need %d" % (n_out, want)
        if use_sdf:
            assert beta is not None and alpha is not None

    @classmethod
    def from_generator(cls, gen) -> "Field":
        """`gen`: a reference `models.generator.Generator` (or anything with the same attributes)."""
        assert not getattr(gen, "use_viewdir", False), \
            "use_viewdir (CARLA, models/generator.py:189-253) is not covered by the fused kernels"
        l0, l2 = gen.decoder.net[0], gen.decoder.net[2]
        for l in (l0, l2):   # lr_multiplier = 1 in the decoder (models/generator.py:295-298)
            assert abs(l.bias_gain - 1.0) < 1e-12 and abs(l.weight_gain * l.in_channels ** 0.5 - 1.0) < 1e-6
        return cls(l0.weight, l0.bias, l2.weight, l2.bias, getattr(gen, "beta", None),
                   getattr(gen, "alpha", None), gen.scene_range, gen.attention_values, gen.use_sdf)

    def tensors(self):
        return self.W0, self.b0, self.W1, self.b1, self.beta, self.alpha

    def to(self, device):
        mv = lambda t: None if t is None else t.to(device)
        return Field(mv(self.W0), mv(self.b0), mv(self.W1), mv(self.b1), mv(self.beta), mv(self.alpha),
                     self.cfg.scene_range, self.cfg.attention_values, self.cfg.use_sdf)


def _as_planes_tensor(xy, xz, yz):
    """The three planes are views `planes[:, k]` of one synthesis output viewed [B,3,32,Hp,Wp]
    (models/generator.py:475-477, 501-503); recover that tensor as [B,96,Hp,Wp] without a copy when
    they are.  The root of the view chain is whatever StyleGAN2 allocated -- for the real
    SynthesisNetwork the grouped-conv / upsample buffer [B*96,1,Hp,Wp] (models/stylegan.py:72-76) --
    so only its storage is trusted, never its shape."""
    base = xy._base
    B, C, Hp, Wp = xy.shape
    per_plane = C * Hp * Wp
    dense = (3 * per_plane, Hp * Wp, Wp, 1)
    esz = xy.element_size()
    if (base is not None and xz._base is base and yz._base is base and base.is_contiguous()
            and base.numel() == 3 * xy.numel() and base.dtype == xy.dtype
            and xy.shape == xz.shape == yz.shape
            and xy.stride() == dense and xz.stride() == dense and yz.stride() == dense
            and xy.data_ptr() == base.data_ptr()
            and xz.data_ptr() == base.data_ptr() + per_plane * esz
            and yz.data_ptr() == base.data_ptr() + 2 * per_plane * esz):
        return base.view(B, 3 * C, Hp, Wp)
    return torch.stack((xy, xz, yz), dim=1).reshape(B, 3 * C, Hp, Wp)


def triplanar_decoder_forward(self, xy, xz, yz, coords, requires_double_backward=False):
    """Replacement for `TriplanarDecoder.forward` (models/generator.py:301-331).

    `coords` [B, N, 1, 3] in box units -> {'features' [B,N,n_out-1], 'density_or_distance' [B,N,1]}.
    The double-backward path (eikonal regulariser, generator.py:306-310, 505-585) runs on
    ops.decoder_double_backward; only a use_viewdir decoder (33 outputs) on that path falls back
    to the saved reference forward."""
    if requires_double_backward:
        # the eikonal regulariser differentiates the output w.r.t. the coordinates with
        # create_graph=True (generator.py:533-536); ops.decoder_double_backward supports exactly that
        l0, l2 = self.net[0], self.net[2]
        if not xy.is_cuda:
            from ._lib import NfiError
            raise NfiError("nfi_b200 ops run on CUDA tensors only (got %s); there is no CPU path" % xy.device)
        if l2.weight.shape[0] <= 16:       # a 33-output (use_viewdir) decoder on this path stays on the reference's PyTorch ops
            assert xy.shape[1] == self.num_input_features == 32
            planes_cl = ops.planes_to_channels_last(_as_planes_tensor(xy, xz, yz))
            B = xy.shape[0]
            y = ops.decoder_double_backward(planes_cl, l0.weight, l0.bias, l2.weight, l2.bias, coords.reshape(B, -1, 3))
            return {"features": y[..., 1:], "density_or_distance": y[..., :1]}
        return self._nfi_reference_forward(xy, xz, yz, coords, requires_double_backward)
    assert xy.shape[1] == self.num_input_features
    assert xz.shape[1] == self.num_input_features
    assert yz.shape[1] == self.num_input_features
    assert self.num_input_features == 32
    planes_cl = ops.planes_to_channels_last(_as_planes_tensor(xy, xz, yz))
    l0, l2 = self.net[0], self.net[2]
    n_out = l2.weight.shape[0]
    cfg = FieldConfig(attention_values=(n_out - 1) if n_out != 4 else 0, use_sdf=False, scene_range=1.0)
    B = xy.shape[0]
    pts = coords.reshape(B, -1, 3)
    y, _ = ops.field_on_points(planes_cl, None, l0.weight, l0.bias, l2.weight, l2.bias, None, None,
                               pts, cfg, mode=1)
    return {"features": y[..., 1:], "density_or_distance": y[..., :1]}


def make_sampler(planes, palette, field: Field, planes_layout: str = "nchw"):
    """Replacement for the `sampler` closure (models/generator.py:587-681) for the
    ['sdf_distance', 'sigma', 'rgb', 'normals', 'semantics', 'coords'] outputs.  `planes`: [B,96,Hp,Wp] / [B,3,32,Hp,Wp] (nchw) or
    [B,3,Hp,Wp,32] (cl)."""
    planes_cl = planes if planes_layout == "cl" else ops.planes_to_channels_last(planes)

    def sampler(x_in, request_sampler_outputs=("sigma", "rgb")):
        for output in request_sampler_outputs:
            assert output in ["sdf_distance", "sigma", "rgb", "normals", "semantics", "coords"]
        bs = x_in.shape[0]
        pts = x_in.reshape(bs, -1, 3)
        out = {}
        want_nrm = "normals" in request_sampler_outputs
        if want_nrm:                            # generator.py:599-602, 614-622: everything detached
            assert field.cfg.use_sdf and torch.is_grad_enabled()
            out["normals"] = ops.sdf_normals_on_points(planes_cl, *field.tensors()[:4], pts, field.cfg)
        if "sdf_distance" in request_sampler_outputs:
            y, _ = ops.field_on_points(planes_cl, None, *field.tensors(), pts, field.cfg, mode=1)
            out["sdf_distance"] = y[..., :1]
        want_sem = "semantics" in request_sampler_outputs
        if want_sem:
            assert field.cfg.attention_values > 0
        if want_sem or "sigma" in request_sampler_outputs or "rgb" in request_sampler_outputs:
            o4, probs = ops.field_on_points(planes_cl, palette, *field.tensors(), pts, field.cfg, mode=0,
                                            want_probs=want_sem)
            if want_nrm:
                o4 = o4.detach()
            if "sigma" in request_sampler_outputs:
                out["sigma"] = o4[..., 0]
            if "rgb" in request_sampler_outputs:
                out["rgb"] = o4[..., 1:]
            if want_sem:
                out["semantics"] = probs
        if "coords" in request_sampler_outputs:
            out["coords"] = x_in
        return out

    return sampler
