"""Test helper: the UNMODIFIED reference `models.generator.Generator` (real StyleGAN2
`SynthesisNetwork`, no stub) and the reference's own `render`, lifted from run.py with `ast`
(oracle/ref_loader.py).  Comes from /root/reference in the build container and from the staged copy
in git-ignored oracle/_ref/ on the GPU box (oracle/build_ref.py)."""
import contextlib
import sys
import types

import torch

import ref_loader


def build_generator(device="cpu", attention_values=10, use_sdf=True, scene_range=1.4, seed=0,
                    sdf_shift=-1.0, use_viewdir=False):
    """Random-init p3d_car-style generator (run.py:560-574 `create_model`).  The SDF bias is shifted so
    that the rendered mask is non-trivial at random init (mean SDF ~ +0.65 otherwise: empty scene; -1.0 gives a mask mean of ~0.37)."""
    m = ref_loader.modules()
    torch.manual_seed(seed)
    G = m.generator.Generator(512, scene_range, attention_values=attention_values, use_viewdir=use_viewdir,
                              use_encoder=False, use_sdf=use_sdf)
    with torch.no_grad():
        G.decoder.net[2].bias[0] += sdf_shift
        if use_viewdir:        # the output layer of the mapper is zero-initialised (generator.py:219-220): constant colours
            G.viewdir_mapper.output.weight.normal_()
            G.viewdir_mapper.output.bias.normal_(std=0.1)
    return G.to(device)


def make_ws(B, device="cpu", seed=1, num_ws=15):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(B, num_ws, 512, generator=g) * 0.5).to(device)


def args_ns(attention_values=10, use_sdf=True, fine_sampling=True, use_viewdir=False):
    return types.SimpleNamespace(use_viewdir=use_viewdir, use_sdf=use_sdf, attention_values=attention_values,
                                 fine_sampling=fine_sampling)


@contextlib.contextmanager
def installed_modules():
    """`nfi_b200.patch.install_modules()` (route (i)) for the duration of a test, then undone."""
    from nfi_b200 import patch
    m = ref_loader.modules()
    keys = ("lib", "lib.nerf_utils", "lib.pose_utils", "models", "models.generator")
    saved = {k: sys.modules.get(k) for k in keys}
    lib_pkg = sys.modules.get("lib")
    saved_attr = getattr(lib_pkg, "nerf_utils", None) if lib_pkg is not None else None
    cls = m.generator.TriplanarDecoder
    fwd = cls.forward
    p2m = m.pose_utils.pose_to_matrix
    try:
        yield patch.install_modules(ref_loader.REF_ROOT)
    finally:
        cls.forward = fwd
        if hasattr(cls, "_nfi_reference_forward"):
            del cls._nfi_reference_forward
        m.pose_utils.pose_to_matrix = p2m
        if hasattr(m.pose_utils, "_nfi_reference_pose_to_matrix"):
            del m.pose_utils._nfi_reference_pose_to_matrix
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        if lib_pkg is not None and saved_attr is not None:
            lib_pkg.nerf_utils = saved_attr
