"""The drop-in installer: module-level swap and the AST swap of `def render`."""
import os
import sys
import textwrap

import pytest

from nfi_b200 import patch

REF = "/root/reference"


def test_ast_swap_replaces_exactly_the_module_level_render(tmp_path):
    script = tmp_path / "toy.py"
    script.write_text(textwrap.dedent('''
        args = "ARGS"
        dataset_config = {"scene_range": 1.0, "white_background": False}
        def helper():
            def render():          # nested: must not be touched
                return "inner"
            return render()
        def render(target_model, height, width):
            return "reference"
        RESULT = (helper(), render)
    '''))
    import nfi_b200.render as nr
    seen = {}
    orig = nr.make_render
    nr.make_render = lambda a, d, zbuffer_depth=False: seen.setdefault("call", (a, d, zbuffer_depth)) and "fused"
    try:
        glb = {"__name__": "toy"}
        exec(patch.compile_with_fused_render(str(script)), glb)
    finally:
        nr.make_render = orig
    assert glb["RESULT"] == ("inner", "fused")
    assert seen["call"] == ("ARGS", glb["dataset_config"], False)


def test_ast_swap_rejects_scripts_without_render(tmp_path):
    script = tmp_path / "none.py"
    script.write_text("x = 1\n")
    with pytest.raises(RuntimeError):
        patch.compile_with_fused_render(str(script))


@pytest.mark.reference
@pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not present")
def test_reference_scripts_have_one_swappable_render():
    for name, zb in (("run.py", False), ("eval_kitti_persp.py", True)):
        path = os.path.join(REF, name)
        if os.path.exists(path):
            code = patch.compile_with_fused_render(path)
            assert code.co_filename == path


@pytest.mark.reference
@pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not present")
def test_install_modules_swaps_nerf_utils_and_decoder():
    saved = {k: sys.modules.get(k) for k in ("lib", "lib.nerf_utils", "lib.pose_utils", "models", "models.generator")}
    saved_path = list(sys.path)
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    sys.dont_write_bytecode = True
    try:
        done = patch.install_modules(REF)
        import nfi_b200.nerf_utils as ours
        assert sys.modules["lib.nerf_utils"] is ours
        assert "lib.nerf_utils" in done
        if "models.generator.TriplanarDecoder.forward" in done:
            gen = sys.modules["models.generator"]
            from nfi_b200 import decoder
            assert gen.TriplanarDecoder.forward is decoder.triplanar_decoder_forward
            assert hasattr(gen.TriplanarDecoder, "_nfi_reference_forward")
            gen.TriplanarDecoder.forward = gen.TriplanarDecoder._nfi_reference_forward
            del gen.TriplanarDecoder._nfi_reference_forward
        if "lib.pose_utils.pose_to_matrix" in done:
            import torch
            pu = sys.modules["lib.pose_utils"]
            assert hasattr(pu, "_nfi_reference_pose_to_matrix")
            # CPU tensors still go to the reference implementation (no CPU path in the library)
            q = torch.nn.functional.normalize(torch.randn(2, 4), dim=-1)
            mat, focal = pu.pose_to_matrix(torch.zeros(2), torch.zeros(2, 2), torch.ones(2), q, True)
            ref = pu._nfi_reference_pose_to_matrix(torch.zeros(2), torch.zeros(2, 2), torch.ones(2), q, True)
            assert torch.equal(mat, ref[0]) and torch.equal(focal, ref[1])
            pu.pose_to_matrix = pu._nfi_reference_pose_to_matrix
            del pu._nfi_reference_pose_to_matrix
    finally:
        sys.path[:] = saved_path
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
