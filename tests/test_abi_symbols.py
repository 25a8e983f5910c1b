"""CPU-side checks of the drop-in boundary: include/nfi_b200.h, the ctypes signature table and the
built library agree on every symbol (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "nfi_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"^\s*(?:int|size_t|const char\s*\*)\s*(nfi_[a-z0-9_]+)\s*\(", src, flags=re.M)))


def test_header_declares_every_bound_symbol():
    from nfi_b200 import _lib
    bound = set(_lib.exported_symbols()) | {"nfi_version", "nfi_last_error"}
    declared = set(_declared())
    assert declared, "no declarations parsed from the header"
    assert declared == bound, (sorted(declared - bound), sorted(bound - declared))


def test_header_cites_the_reference():
    src = open(HEADER).read()
    for cite in ("lib/nerf_utils.py", "models/generator.py", "run.py"):
        assert cite in src


def test_library_loads_and_exports_every_symbol():
    from nfi_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), "libnfi_b200.so does not export %s" % name
    lib.nfi_version.restype = ctypes.c_int
    assert lib.nfi_version() > 0
    lib.nfi_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.nfi_last_error(), bytes)


def test_bad_arguments_are_reported_not_fatal():
    """NULL pointers must come back as NFI_ERR_BAD_ARG with a message, before any CUDA call."""
    from nfi_b200 import _lib
    lib = _lib.lib()
    rc = lib.nfi_stratified(None, None, None, 4, 8, 0, 0, 0, None, None)
    assert rc < 0
    assert b"nfi_stratified" in lib.nfi_last_error()


def test_product_path_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nfi_b200 import nerf_utils
    with pytest.raises((RuntimeError, AssertionError, ValueError)):
        nerf_utils.compute_near_far_planes(torch.zeros(1, 2, 2, 3), torch.ones(1, 2, 2, 3), 1.4)
