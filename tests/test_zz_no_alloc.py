"""skp_attn_map_fwd_static (no caller workspace, no allocation: csrc/attn_noalloc.cu) against the existing
skp_attn_map_fwd, through the C ABI, on two calls in a row that share the library's static scratch array.

Collected last on purpose (pytest -x at the end of a round): a failure here must not hide the other tests.
Passes alone in a fresh process and after the other tests (it only depends on the library).
Tolerance = the forward's parity bar used everywhere in tests/ (rtol 1e-3, atol 1e-7); the two entry points run the
same kernels on the same operands, so the results are in fact expected to be identical.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _layers(spec, T, B, seed):
    g = torch.Generator().manual_seed(seed)
    qs = [(torch.randn(B, s * s, c, generator=g) * 1.5).to(DEV) for s, c, _ in spec]
    ks = [torch.randn(1, T, c, generator=g).to(DEV) for _, c, _ in spec]
    return qs, ks


def _array(_lib, spec, qs, ks, B):
    arr = (_lib.AttnLayer * len(spec))()
    for i, (s, c, h) in enumerate(spec):
        arr[i].q, arr[i].k = qs[i].data_ptr(), ks[i].data_ptr()
        arr[i].s, arr[i].C, arr[i].heads = s, c, h
        arr[i].k_shared = int(B > 1)
        arr[i].scale, arr[i].weight = (c // h) ** -0.5, 1.0 / len(spec)
    return arr


def _both(_lib, spec, T, R, B, seed, token_idx=None):
    """(maps, stats) of skp_attn_map_fwd with a torch-allocated workspace and of skp_attn_map_fwd_static."""
    lib = _lib.load()
    qs, ks = _layers(spec, T, B, seed)
    arr = _array(_lib, spec, qs, ks, B)
    n = len(spec)
    prec = _lib.PREC_TC_BF16X3            # the tcgen05 path is the one that uses scratch
    t_out = T if token_idx is None else token_idx.numel()
    idx_p = None if token_idx is None else C.c_void_p(token_idx.data_ptr())
    n_idx = 0 if token_idx is None else token_idx.numel()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    heads = sum(h for _, _, h in spec)
    out = []
    need = lib.skp_attn_map_workspace_bytes_for(arr, n, B, T, R, prec, _lib.WS_FWD)
    assert 0 < need <= (96 << 20)
    ws = torch.empty(need, dtype=torch.uint8, device=DEV)
    for static in (False, True):
        maps = torch.full((B, t_out, R, R), float("nan"), device=DEV)
        stats = torch.full((B, heads, R * R, 2), float("nan"), device=DEV)
        if static:
            rc = lib.skp_attn_map_fwd_static(arr, n, B, T, R, idx_p, n_idx, 0, C.c_void_p(maps.data_ptr()),
                                             C.c_void_p(stats.data_ptr()), prec, stream)
        else:
            rc = lib.skp_attn_map_fwd(arr, n, B, T, R, idx_p, n_idx, 0, C.c_void_p(maps.data_ptr()),
                                      C.c_void_p(stats.data_ptr()), prec, C.c_void_p(ws.data_ptr()), need, stream)
        assert rc == 0, lib.skp_last_error().decode()
        torch.cuda.synchronize()
        out.append((maps, stats))
    return out


def test_static_scratch_forward_matches_workspace_forward_twice():
    from stablekeypoints_b200 import _lib
    # call 1: two source sizes (64- and 128-tap windows), ragged T, all tokens; call 2: a different, smaller problem
    # with selected tokens, so stale contents of the shared array from call 1 would show
    cases = [([(16, 128, 2), (32, 64, 1)], 70, 128, 2, None),
             ([(4, 64, 4), (4, 64, 4), (8, 32, 2)], 37, 16, 1, torch.tensor([5, 0, 36, 17], device=DEV))]
    for seed, (spec, T, R, B, idx) in enumerate(cases):
        (m0, s0), (m1, s1) = _both(_lib, spec, T, R, B, 20 + seed, idx)
        assert torch.isfinite(m1).all() and torch.isfinite(s1).all()
        assert torch.allclose(m1, m0, rtol=1e-3, atol=1e-7), float((m1 - m0).abs().max())
        assert torch.allclose(s1, s0, rtol=1e-3, atol=1e-6)
        if idx is None:
            assert torch.allclose(m1.sum(1), torch.ones(B, R, R, device=DEV), atol=2e-5)


def test_static_scratch_rejects_problems_beyond_the_limit():
    from stablekeypoints_b200 import _lib
    lib = _lib.load()
    spec = [(16, 1280, 20)] * 3 + [(32, 640, 10)]          # BASELINE capture set, 2 images: 138.7 MiB > 96 MiB
    arr = (_lib.AttnLayer * len(spec))()
    for i, (s, c, h) in enumerate(spec):
        arr[i].s, arr[i].C, arr[i].heads, arr[i].k_shared = s, c, h, 1
        arr[i].scale, arr[i].weight = (c // h) ** -0.5, 1.0 / len(spec)
    dummy = torch.empty(16, device=DEV)
    rc = lib.skp_attn_map_fwd_static(arr, len(spec), 2, 500, 128, None, 0, 0, C.c_void_p(dummy.data_ptr()), None,
                                     _lib.PREC_TC_BF16X3, None)
    assert rc == -4                                          # SKP_ERR_WORKSPACE, before any launch
    assert b"static array holds" in lib.skp_last_error()
