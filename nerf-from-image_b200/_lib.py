"""ctypes binding of libnfi_b200.so (include/nfi_b200.h) -- the only way the package reaches the GPU.

There is no CPU path and no PyTorch fallback: if the library is missing (not built) the import
of any op raises, and every op raises on a non-CUDA tensor.  Pointers are raw device addresses
(`tensor.data_ptr()`), the stream is torch's *current* stream of the tensor's device, and ctypes
releases the GIL for the duration of each call (safe under nn.DataParallel worker threads).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
# NFI_B200_LIB: developer switch to load another build of the same library (kernel A/B runs)
LIB_PATH = os.environ.get("NFI_B200_LIB") or os.path.join(HERE, "lib", "libnfi_b200.so")


class NfiError(RuntimeError):
    pass


class NfiField(Structure):
    """Mirror of `struct NfiField` in include/nfi_b200.h."""
    _fields_ = [("W0", c_void_p), ("b0", c_void_p), ("W1", c_void_p), ("b1", c_void_p),
                ("beta", c_void_p), ("alpha", c_void_p), ("n_out", c_int32),
                ("attention_values", c_int32), ("use_sdf", c_int32), ("scene_range", c_float),
                ("planes_bf16", c_int32)]


class NfiRender(Structure):
    """Mirror of `struct NfiRender` in include/nfi_b200.h."""
    _fields_ = [("field", NfiField), ("planes_cl", c_void_p), ("palette", c_void_p), ("cam2world", c_void_p),
                ("focal", c_void_p), ("center", c_void_p), ("bbox", c_void_p), ("u_strat", c_void_p),
                ("u_pdf", c_void_p), ("seed", c_uint64), ("offset", c_uint64), ("B", c_int32), ("H", c_int32),
                ("W", c_int32), ("S", c_int32), ("Hp", c_int32), ("Wp", c_int32), ("focal_stride", c_int32),
                ("randomize", c_int32), ("white_background", c_int32), ("zbuffer_depth", c_int32)]


_P = c_void_p
_SIGS = {
    "nfi_pose_to_matrix_fwd": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, _P, _P, _P]),
    "nfi_pose_to_matrix_bwd": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, _P, _P, _P, _P, _P, _P, _P]),
    "nfi_render_workspace_bytes": (ctypes.c_size_t, [POINTER(NfiRender), c_int]),
    "nfi_render_fwd": (c_int, [POINTER(NfiRender), _P, _P, _P, _P, _P]),
    "nfi_render_bwd": (c_int, [POINTER(NfiRender), _P, _P, _P, _P, _P, _P, _P, _P, c_int, _P]),
    "nfi_version": (c_int, []),
    "nfi_last_error": (c_char_p, []),
    "nfi_ray_bundle_fwd": (c_int, [_P, _P, c_int, _P, _P, c_int, c_int, c_int, c_int, _P, _P, _P]),
    "nfi_ray_bundle_bwd": (c_int, [_P, _P, c_int, _P, _P, c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P]),
    "nfi_near_far": (c_int, [_P, _P, c_int64, c_float, _P, _P, _P, _P]),
    "nfi_near_far_static": (c_int, [_P, _P, c_int64, c_float, _P, _P, _P, _P]),
    "nfi_stratified": (c_int, [_P, _P, _P, c_int64, c_int, c_int, c_uint64, c_uint64, _P, _P]),
    "nfi_points_fwd": (c_int, [_P, _P, _P, c_int64, c_int, _P, _P]),
    "nfi_points_bwd": (c_int, [_P, _P, c_int64, c_int, _P, _P, _P]),
    "nfi_planes_to_cl": (c_int, [_P, c_int, c_int, c_int, _P, _P]),
    "nfi_planes_grad_from_cl": (c_int, [_P, c_int, c_int, c_int, _P, _P]),
    "nfi_field_fwd": (c_int, [POINTER(NfiField), _P, _P, c_int, c_int, c_int, _P, _P, _P, _P, c_int64,
                              c_int, c_int, c_int, c_int, _P, _P, _P]),
    "nfi_field_bwd": (c_int, [POINTER(NfiField), _P, _P, c_int, c_int, c_int, _P, _P, _P, _P, c_int64,
                              c_int, c_int, c_int, _P, _P, _P, _P, _P, c_int, _P]),
    "nfi_field_bwd_train_workspace_bytes": (ctypes.c_size_t, [c_int64]),
    "nfi_field_bwd_train": (c_int, [POINTER(NfiField), _P, _P, c_int, c_int, c_int, _P, _P, _P, c_int64, c_int, c_int,
                                    _P, _P, _P, _P, _P, _P, ctypes.c_size_t, _P]),
    "nfi_decoder_gradx_fwd": (c_int, [POINTER(NfiField), _P, c_int, c_int, c_int, _P, _P, c_int64, _P, _P]),
    "nfi_decoder_gradx_bwd": (c_int, [POINTER(NfiField), _P, c_int, c_int, c_int, _P, _P, _P, c_int64, _P, _P, _P, _P, _P]),
    "nfi_decoder_mlp_fwd": (c_int, [POINTER(NfiField), _P, c_int64, _P, _P, _P]),
    "nfi_weights_only": (c_int, [_P, c_int, _P, _P, c_int64, c_int, c_int, _P, _P]),
    "nfi_sample_pdf": (c_int, [_P, _P, c_int64, c_int, c_int, _P, _P, _P, _P]),
    "nfi_importance": (c_int, [_P, c_int, _P, _P, c_int64, c_int, _P, _P, _P]),
    "nfi_cumprod_exclusive": (c_int, [_P, c_int64, c_int, _P, _P]),
    "nfi_composite_fwd": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int64, c_int, c_int, _P, _P, _P, _P, _P, _P]),
    "nfi_composite_bwd": (c_int, [_P, _P, _P, _P, _P, c_int, c_int64, c_int, c_int, _P, _P, _P, _P, _P, _P, _P, _P]),
    "nfi_merge_composite_fwd": (c_int, [_P, _P, _P, _P, _P, c_int64, c_int, c_int, _P, c_int64, _P, _P, _P, _P, _P, _P]),
    "nfi_merge_composite_bwd": (c_int, [_P, _P, _P, _P, _P, c_int64, c_int, c_int, _P, _P, _P, _P, _P]),
}

_lib = None


def exported_symbols():
    """Names include/nfi_b200.h declares (used by the CPU-side symbol test)."""
    return sorted(_SIGS)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NfiError("libnfi_b200.so is not built (%s): run `python __graft_entry__.py build` "
                           "-- there is no CPU fallback" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            f = getattr(l, name)      # AttributeError if the header and the library disagree
            f.restype = res
            f.argtypes = args
        _lib = l
    return _lib


def ptr(t):
    """Device pointer of an fp32/int32/uint8 contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise NfiError("nfi_b200 ops run on CUDA tensors only (got %s); there is no CPU path" % t.device)
    if not t.is_contiguous():
        raise NfiError("internal: non-contiguous tensor passed to the C ABI")
    return c_void_p(t.data_ptr())


def f32c(t):
    """fp32 contiguous view/copy (None passes through)."""
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def stream_of(t):
    return c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


# kernels launched per C-ABI call (everything else launches exactly one)
KERNELS_PER_CALL = {"nfi_near_far": 2, "nfi_near_far_static": 2, "nfi_field_bwd_train": 2}


class LaunchProfile:
    """Optional accounting used by bench.py: counts this library's kernel launches and brackets
    chosen entry points with CUDA events on the launching stream.  Off by default (zero cost)."""

    def __init__(self):
        self.enabled = False
        self.launches = 0
        self.timed = ()
        self.events = {}

    def start(self, timed=()):
        self.enabled, self.launches, self.timed, self.events = True, 0, tuple(timed), {}

    def stop(self):
        self.enabled = False

    def durations_ms(self):
        """{entry point: [ms per call]} -- call after a synchronize."""
        return {k: [a.elapsed_time(b) for a, b in v] for k, v in self.events.items()}


PROFILE = LaunchProfile()
LAUNCHES = [0]            # kernels of this library launched so far by this process (always counted)


def call(name, device_tensor, *args):
    """Run `name(*args, stream)` with `device_tensor`'s device current; raise on a negative code."""
    l = lib()
    with torch.cuda.device(device_tensor.device):
        ev = None
        LAUNCHES[0] += KERNELS_PER_CALL.get(name, 1)
        if PROFILE.enabled:
            PROFILE.launches += KERNELS_PER_CALL.get(name, 1)
            if name in PROFILE.timed and not torch.cuda.is_current_stream_capturing():
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                ev[0].record()
        rc = getattr(l, name)(*args, stream_of(device_tensor))
        if ev is not None:
            ev[1].record()
            PROFILE.events.setdefault(name, []).append(ev)
    if rc != 0:
        msg = l.nfi_last_error().decode("utf-8", "replace")
        raise NfiError("%s failed (%d): %s" % (name, rc, msg))


def make_field_struct(W0, b0, W1, b1, beta, alpha, attention_values, use_sdf, scene_range, planes_bf16=False):
    """Build an NfiField; the returned `keep` list must outlive the call (holds the tensors)."""
    keep = [f32c(W0), f32c(b0), f32c(W1), f32c(b1),
            f32c(beta) if use_sdf else None, f32c(alpha) if use_sdf else None]
    n_out = int(keep[2].shape[0])
    assert keep[0].shape == (64, 32), "decoder.net[0].weight must be [64, 32] (models/generator.py:294-296)"
    assert keep[2].shape[1] == 64
    f = NfiField(ptr(keep[0]), ptr(keep[1]), ptr(keep[2]), ptr(keep[3]), ptr(keep[4]), ptr(keep[5]),
                 n_out, int(attention_values), 1 if use_sdf else 0, float(scene_range), 1 if planes_bf16 else 0)
    return f, keep
