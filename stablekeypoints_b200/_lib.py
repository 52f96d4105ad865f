"""ctypes binding of libskp_b200.so (the C ABI declared in include/skp_b200.h).

The library is the product path; there is no CPU or eager-PyTorch fallback.  Importing this
module without the built library raises, and every wrapper raises on non-CUDA tensors.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libskp_b200.so")

PREC_AUTO, PREC_SIMT_F32, PREC_TC_BF16, PREC_TC_BF16X3 = 0, 1, 2, 3
WS_ANY, WS_FWD, WS_BWD_STATS, WS_BWD_NOSTATS = 0, 1, 2, 3
PRECISIONS = {"auto": PREC_AUTO, "simt_f32": PREC_SIMT_F32, "tc_bf16": PREC_TC_BF16, "tc_bf16x3": PREC_TC_BF16X3}


class AttnLayer(C.Structure):
    """struct skp_attn_layer"""
    _fields_ = [("q", C.c_void_p), ("k", C.c_void_p), ("dq", C.c_void_p), ("dk", C.c_void_p),
                ("s", C.c_int32), ("C", C.c_int32), ("heads", C.c_int32), ("k_shared", C.c_int32),
                ("scale", C.c_float), ("weight", C.c_float)]


# name -> (restype, argtypes); must list every symbol include/skp_b200.h declares
_P, _I, _F, _Z, _L = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int64
SIGNATURES = {
    "skp_version": (_I, []),
    "skp_last_error": (C.c_char_p, []),
    "skp_device_arch": (_I, [_I]),
    "skp_kernel_launches": (C.c_longlong, []),
    "skp_attn_map_workspace_bytes": (_Z, [C.POINTER(AttnLayer), _I, _I, _I, _I, _I]),
    "skp_attn_map_workspace_bytes_for": (_Z, [C.POINTER(AttnLayer), _I, _I, _I, _I, _I, _I]),
    "skp_attn_map_fwd": (_I, [C.POINTER(AttnLayer), _I, _I, _I, _I, _P, _I, _L, _P, _P, _I, _P, _Z, _P]),
    "skp_attn_map_fwd_static": (_I, [C.POINTER(AttnLayer), _I, _I, _I, _I, _P, _I, _L, _P, _P, _I, _P]),
    "skp_attn_map_bwd_reuse": (_I, [C.POINTER(AttnLayer), _I, _I, _I, _I, _P, _I, _L, _P, _P, _I, _P, _Z, _P, _Z, _P]),
    "skp_attn_map_bwd": (_I, [C.POINTER(AttnLayer), _I, _I, _I, _I, _P, _I, _L, _P, _P, _I, _P, _Z, _P]),
    "skp_bilinear_upsample": (_I, [_P, _I, _I, _I, _P, _I, _I, _P]),
    "skp_bicubic_upsample_nhwc": (_I, [_P, _I, _I, _I, _P, _I, _P]),
    "skp_affine_warp": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "skp_affine_warp_bwd": (_I, [_P, _P, _I, _I, _I, _I, _P, _P]),
    "skp_plane_argmax": (_I, [_P, _I, _I, _I, _P, _P, _P]),
    "skp_find_k_max_pixels": (_I, [_P, _I, _I, _I, _I, _P, _P]),
    "skp_weighted_avg_readout": (_I, [_P, _I, _I, _I, _F, _I, _P, _P]),
    "skp_plane_stats": (_I, [_P, _I, _I, _I, _F, _F, _I, _P, _P, _P, _P]),
    "skp_gaussian_circles": (_I, [_P, _I, _I, _I, _F, _P, _P]),
    "skp_rank_ascending": (_I, [_P, _I, _I, _P, _P]),
    "skp_rank_ascending_batched": (_I, [_P, _I, _I, _I, _P, _P]),
    "skp_furthest_point_sampling_batched": (_I, [_P, _I, _I, _I, _I, _P, _I, _I, _P, _P, _P]),
    "skp_keypoint_losses_batched_fwd": (_I, [_P, _P, _P, _L, _I, _I, _I, _I, _I, _P, _F, _I, _P, _P]),
    "skp_keypoint_losses_batched_bwd": (_I, [_P, _P, _P, _L, _I, _I, _I, _I, _I, _P, _F, _I, _P, _P, _P, _P]),
    "skp_furthest_point_sampling": (_I, [_P, _I, _I, _I, _P, _I, _I, _P, _P, _P]),
    "skp_keypoint_losses_fwd": (_I, [_P, _P, _P, _I, _I, _I, _P, _F, _I, _P, _P]),
    "skp_keypoint_losses_bwd": (_I, [_P, _P, _P, _I, _I, _I, _P, _F, _I, _F, _F, _P, _P, _P]),
    "skp_tta_accumulate": (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _I, _I, _P]),
    "skp_tta_finalize": (_I, [_P, _P, _L, _P, _P]),
}

_lib = None
launches = 0   # kernels-launching C-ABI calls made through this binding (bench's gpu_launches evidence)


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (or ./build.sh). "
                "There is no CPU fallback for the stablekeypoints_b200 hot path.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError here == header and library out of sync
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {load().skp_last_error().decode()}")


# Optional per-entry-point device timing (bench.py's live roofline numbers): when `timing` is a
# dict, every call is bracketed by CUDA events on the stream it launches on (the last argument
# of every entry point is the stream), and `timing_summary()` turns them into ms after a sync.
timing = None


def timing_start() -> None:
    global timing
    timing = {}


def timing_summary() -> dict:
    """{entry point: {"calls": n, "ms": total device ms}}; stops timing."""
    global timing
    out = {}
    if timing:
        torch.cuda.synchronize()
        for name, evs in timing.items():
            out[name] = {"calls": len(evs), "ms": sum(a.elapsed_time(b) for a, b in evs)}
    timing = None
    return out


def call(name: str, *args) -> None:
    global launches
    launches += 1
    if timing is None:
        check(getattr(load(), name)(*args), name)
        return
    st = torch.cuda.ExternalStream(args[-1].value) if args[-1].value else torch.cuda.default_stream()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    check(getattr(load(), name)(*args), name)
    b.record(st)
    timing.setdefault(name, []).append((a, b))


def ptr(t):
    """device pointer of a CUDA tensor (None -> NULL); refuses CPU tensors: no CPU path exists."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("stablekeypoints_b200 kernels need CUDA tensors (there is no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("stablekeypoints_b200 kernels need contiguous tensors")
    return C.c_void_p(t.data_ptr())


def stream_of(t) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)
