"""Loader for the UNMODIFIED reference (yuliangguo/nerf-from-image) -- TEST INFRASTRUCTURE ONLY.

Nothing in the product package may import this module.  It exists so that
`oracle/make_golden.py` and `tests/test_oracle_vs_reference.py` can run the
reference's own code (in the build container, where /root/reference is mounted)
and pin `oracle/nfi_oracle.py` + the golden fixtures under tests/golden/ to it.
On the GPU box /root/reference does not exist; the loader then uses the unmodified files staged in
git-ignored `oracle/_ref/` by `oracle/build_ref.py` (they travel with the snapshot), and
`available()` is False only when neither exists.

`run.py` cannot be imported (argparse + missing deps at module level,
run.py:42), so `render` (run.py:176-350) and the z-buffer variant
(eval_kitti_persp.py:43-230) are lifted with `ast` and exec'd in a small
namespace -- the reference source is never copied into this repository.
"""
import ast
import os
import sys
import types

_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _pick_root():
    """The mounted checkout where it exists (build container), else the copy staged by
    oracle/build_ref.py into git-ignored oracle/_ref/ (this is what the GPU box sees)."""
    cands = [os.environ.get("NFI_REFERENCE"), "/root/reference", _STAGED]
    for c in cands:
        if c and os.path.isfile(os.path.join(c, "lib", "nerf_utils.py")):
            return c
    return "/root/reference"


REF_ROOT = _pick_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "lib", "nerf_utils.py"))


_mods = None


def modules():
    """Import lib.nerf_utils, lib.ops, lib.pose_utils, models.generator from the reference."""
    global _mods
    if _mods is None:
        if not available():
            raise RuntimeError("reference checkout not present at %s" % REF_ROOT)
        sys.dont_write_bytecode = True
        if REF_ROOT not in sys.path:
            sys.path.insert(0, REF_ROOT)
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            from lib import nerf_utils, ops, pose_utils  # noqa
            from models import generator, stylegan  # noqa
        _mods = types.SimpleNamespace(nerf_utils=nerf_utils, ops=ops, pose_utils=pose_utils,
                                      generator=generator, stylegan=stylegan)
    return _mods


def lift_render(script="run.py", *, use_viewdir=False, use_sdf=True, attention_values=10,
                fine_sampling=True, scene_range=1.4, white_background=False, nerf_utils=None):
    """Return the reference's `render` function bound to the given args/dataset_config globals.
    `nerf_utils`: the module the lifted function sees under that name (default: the reference's own
    lib.nerf_utils; the drop-in tests pass the replacement module, as `import lib.nerf_utils` would
    resolve after `nfi_b200.patch.install_modules()`)."""
    import torch
    import torch.nn.functional as F
    m = modules()
    src = open(os.path.join(REF_ROOT, script)).read()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "render"]
    assert len(fn) == 1
    mod = ast.Module(body=fn, type_ignores=[])
    ns = {
        "torch": torch, "F": F, "nerf_utils": nerf_utils or m.nerf_utils, "pose_utils": m.pose_utils,
        "args": types.SimpleNamespace(use_viewdir=use_viewdir, use_sdf=use_sdf,
                                      attention_values=attention_values,
                                      fine_sampling=fine_sampling),
        "dataset_config": {"scene_range": scene_range, "white_background": white_background},
    }
    exec(compile(mod, os.path.join(REF_ROOT, script), "exec"), ns)
    return ns["render"], ns
