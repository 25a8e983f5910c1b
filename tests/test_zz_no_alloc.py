"""`nfi_near_far_static` -- the allocation-free variant of A2 (include/nfi_b200.h; no counterpart in the
reference): called through the C ABI and compared with
the existing entry point `nfi_near_far` on two calls in a row that share the library's static scratch.
Bit-exact, like every other A2 test (the two entry points run the same kernels).

Collected last on purpose (file name).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_RAYS = 1 << 24      # NFI_NEARFAR_STATIC_MAX_RAYS


def _rays(B, H, W, seed, scene_range):
    from nfi_b200 import ops, synth
    cam, focal = synth.make_cameras(B, seed=seed, perspective=True)
    o, d = ops.ray_bundle(H, W, focal.cuda(), cam.cuda(), None, None, normalize=True)
    return o.reshape(-1, 3).contiguous(), d.reshape(-1, 3).contiguous()


def _both(lib, o, d, scene_range):
    """-> ((near, far, hits) of nfi_near_far, the same of nfi_near_far_static), raw ctypes calls."""
    R = o.shape[0]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    n0, f0 = torch.empty(R, device="cuda"), torch.empty(R, device="cuda")
    ws = torch.zeros(16384, dtype=torch.uint8, device="cuda")
    rc = lib.nfi_near_far(P(o), P(d), R, scene_range, P(n0), P(f0), P(ws), stream)
    assert rc == 0, lib.nfi_last_error()
    n1, f1 = torch.full((R,), -7.0, device="cuda"), torch.full((R,), -7.0, device="cuda")
    hits = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.nfi_near_far_static(P(o), P(d), R, scene_range, P(n1), P(f1), P(hits), stream)
    assert rc == 0, lib.nfi_last_error()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return (n0, f0, int(ws[:4].view(torch.int32).item())), (n1, f1, int(hits.item())), free_before - free_after


def _same(a, b):
    return bool(((a == b) | (a.isnan() & b.isnan())).all())


def test_near_far_static_matches_near_far_on_two_calls_in_a_row():
    from nfi_b200 import _lib
    lib = _lib.lib()
    # call 1: many rays (all 1024 partial slots of the scratch in use), small box -> some rays miss
    o, d = _rays(4, 256, 320, seed=2, scene_range=0.7)
    ref, got, _ = _both(lib, o, d, 0.7)
    assert 0 < ref[2] < o.shape[0], "fixture: want hits and misses"
    assert got[2] == ref[2] and _same(got[0], ref[0]) and _same(got[1], ref[1])
    # call 2, right after, on the SAME static scratch: fewer rays (a handful of partial slots; the rest
    # still holds call 1's values and must not be read), other cameras, other box
    o, d = _rays(1, 24, 40, seed=5, scene_range=1.4)
    ref, got, grew = _both(lib, o, d, 1.4)
    assert ref[2] > 0
    assert got[2] == ref[2] and _same(got[0], ref[0]) and _same(got[1], ref[1])
    assert grew == 0, "device memory in use changed across an allocation-free call (%d bytes)" % grew
    # hits_out is optional
    R = o.shape[0]
    n, f = torch.empty(R, device="cuda"), torch.empty(R, device="cuda")
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.nfi_near_far_static(P(o), P(d), R, 1.4, P(n), P(f), None,
                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert rc == 0 and _same(n, ref[0]) and _same(f, ref[1])


def test_near_far_static_rejects_sizes_beyond_its_maximum():
    from nfi_b200 import _lib
    lib = _lib.lib()
    t = torch.zeros(8, 3, device="cuda")
    P = lambda x: ctypes.c_void_p(x.data_ptr())
    rc = lib.nfi_near_far_static(P(t), P(t), MAX_RAYS + 1, 1.0, P(t), P(t), None, None)   # rejected before any launch
    assert rc == -1                                                                      # NFI_ERR_BAD_ARG
    msg = lib.nfi_last_error().decode()
    assert "NFI_NEARFAR_STATIC_MAX_RAYS" in msg and str(MAX_RAYS + 1) in msg
    assert lib.nfi_near_far_static(None, P(t), 8, 1.0, P(t), P(t), None, None) == -1
    torch.cuda.synchronize()
