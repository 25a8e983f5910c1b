"""Install the drop-ins into an unmodified reference checkout (SURVEY.md 8(b) "how the swap is
installed").  Nothing under the reference tree is edited.

Route (i), stage level -- `install_modules()`:
    sys.modules['lib.nerf_utils'] <- nfi_b200.nerf_utils            (lib/nerf_utils.py, 7 functions)
    models.generator.TriplanarDecoder.forward <- decoder.triplanar_decoder_forward
                                                                    (models/generator.py:301-331)
    lib.pose_utils.pose_to_matrix <- nfi_b200.pose_utils.pose_to_matrix for CUDA tensors
                                                                    (lib/pose_utils.py:48-75)
Route (ii), fused -- `python -m nfi_b200.patch <script.py> [args...]` (or `run_script`): the script
is compiled with its module-level `def render` (run.py:176-350; a plain global that
ParallelModel.forward looks up at call time, run.py:597) replaced by `nfi_b200.render.make_render`
bound to the script's own `args` / `dataset_config` globals.
"""
import ast
import importlib
import os
import runpy
import sys
import types


def install_modules(reference_root=None, patch_decoder=True):
    """Route (i).  Call before the reference script imports `lib.nerf_utils`.  Returns the names
    that were replaced."""
    from . import decoder as _decoder
    from . import nerf_utils as _nu

    if reference_root and reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    done = []
    lib_pkg = sys.modules.get("lib")
    if lib_pkg is None:
        try:
            lib_pkg = importlib.import_module("lib")
        except ImportError:
            lib_pkg = types.ModuleType("lib")
            lib_pkg.__path__ = []
            sys.modules["lib"] = lib_pkg
    sys.modules["lib.nerf_utils"] = _nu
    setattr(lib_pkg, "nerf_utils", _nu)
    done.append("lib.nerf_utils")
    # N3: lib.pose_utils keeps everything but pose_to_matrix (the rest is off the hot path)
    try:
        pu = importlib.import_module("lib.pose_utils")
    except ImportError:
        pu = None
    if pu is not None and hasattr(pu, "pose_to_matrix"):
        from . import pose_utils as _pu
        if not hasattr(pu, "_nfi_reference_pose_to_matrix"):
            pu._nfi_reference_pose_to_matrix = pu.pose_to_matrix

            def pose_to_matrix(z0, t2, s, q, camera_flipped):
                if not q.is_cuda:                      # CPU callers (dataset preparation) keep the original
                    return pu._nfi_reference_pose_to_matrix(z0, t2, s, q, camera_flipped)
                return _pu.pose_to_matrix(z0, t2, s, q, camera_flipped)

            pu.pose_to_matrix = pose_to_matrix
        done.append("lib.pose_utils.pose_to_matrix")
    if patch_decoder:
        try:
            gen = importlib.import_module("models.generator")
        except ImportError:
            gen = None
        if gen is not None and hasattr(gen, "TriplanarDecoder"):
            cls = gen.TriplanarDecoder
            if not hasattr(cls, "_nfi_reference_forward"):
                cls._nfi_reference_forward = cls.forward     # double-backward path keeps using it
                cls.forward = _decoder.triplanar_decoder_forward
            done.append("models.generator.TriplanarDecoder.forward")
    return done


class _RenderSwap(ast.NodeTransformer):
    """Replace the module-level `def render(...)` by an assignment evaluated at that point of the
    script, so it sees the script's `args` and `dataset_config` globals like the original."""

    def __init__(self, zbuffer_depth):
        self.zbuffer_depth = zbuffer_depth
        self.swapped = 0

    def visit_Module(self, node):
        body = []
        for stmt in node.body:
            if isinstance(stmt, ast.FunctionDef) and stmt.name == "render":
                src = ("render = __import__('nfi_b200.render', fromlist=['make_render'])"
                       ".make_render(args, dataset_config, zbuffer_depth=%r)" % self.zbuffer_depth)
                new = ast.parse(src).body[0]
                body.append(ast.copy_location(new, stmt))
                self.swapped += 1
            else:
                body.append(stmt)
        node.body = body
        return node


def compile_with_fused_render(path, zbuffer_depth=None):
    """Source of `path` -> code object with `render` swapped.  `zbuffer_depth` defaults to True for
    the perspective-eval scripts (their render converts depth, eval_kitti_persp.py:221-228)."""
    with open(path) as f:
        tree = ast.parse(f.read(), filename=path)
    if zbuffer_depth is None:
        zbuffer_depth = "_persp" in os.path.basename(path)
    swap = _RenderSwap(zbuffer_depth)
    tree = ast.fix_missing_locations(swap.visit(tree))
    if swap.swapped != 1:
        raise RuntimeError("%s: expected exactly one module-level `def render`, found %d" % (path, swap.swapped))
    return compile(tree, path, "exec")


def run_script(path, argv=(), stage_level_only=False):
    """Route (ii): run a reference script as __main__ with the fused render installed."""
    path = os.path.abspath(path)
    root = os.path.dirname(path)
    if root not in sys.path:
        sys.path.insert(0, root)
    install_modules(root)
    sys.argv = [path] + list(argv)
    if stage_level_only:
        return runpy.run_path(path, run_name="__main__")
    code = compile_with_fused_render(path)
    glb = {"__name__": "__main__", "__file__": path, "__builtins__": __builtins__}
    exec(code, glb)
    return glb


if __name__ == "__main__":
    if len(sys.argv) < 2:
        raise SystemExit("usage: python -m nfi_b200.patch <reference script.py> [script args...]")
    run_script(sys.argv[1], sys.argv[2:])
