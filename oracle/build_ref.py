"""Recipe for `oracle/_ref/` -- TEST / BASELINE INFRASTRUCTURE ONLY (nothing in the product package
may import it).

The reference (yuliangguo/nerf-from-image) is pure Python: the hot path and the generator that feeds
it import on torch + numpy alone.  `/root/reference` does not exist on the GPU box, so this script
stages the few UNMODIFIED reference files the drop-in tests and the reference arm of `bench.py` need
into `oracle/_ref/` -- a directory that is git-ignored (reference sources never enter this
repository's history) but NOT gpurun-ignored, so it travels to the GPU box like a built `.so`.

    python oracle/build_ref.py            # copy (only where /root/reference is mounted)

`__graft_entry__.build()` runs it when the reference checkout is present.  Files staged:
  lib/nerf_utils.py, lib/ops.py, lib/pose_utils.py        (the seven hot-path functions, grid_sample2d)
  models/generator.py, models/stylegan.py                  (Generator, TriplanarDecoder, SynthesisNetwork)
  run.py, eval_kitti_persp.py                              (`def render` is lifted from them with `ast`;
                                                            the scripts themselves are never executed)
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SRC = os.environ.get("NFI_REFERENCE", "/root/reference")
FILES = ["lib/nerf_utils.py", "lib/ops.py", "lib/pose_utils.py", "models/generator.py", "models/stylegan.py",
         "run.py", "eval_kitti_persp.py", "LICENSE"]


def build(verbose=False):
    """-> DEST when staged (or already there), None when neither the checkout nor a staged copy exists."""
    if not os.path.isfile(os.path.join(SRC, "lib", "nerf_utils.py")):
        return DEST if os.path.isfile(os.path.join(DEST, "lib", "nerf_utils.py")) else None
    for rel in FILES:
        s, d = os.path.join(SRC, rel), os.path.join(DEST, rel)
        if not os.path.exists(s):
            continue
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if not (os.path.exists(d) and filecmp.cmp(s, d, shallow=False)):
            shutil.copyfile(s, d)
            if verbose:
                print("staged", rel, file=sys.stderr)
    return DEST


if __name__ == "__main__":
    print(build(verbose=True))
