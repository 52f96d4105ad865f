"""Tensor-level wrappers over the C ABI (one function per entry point of include/skp_b200.h),
plus the torch.autograd.Function glue.  PyTorch is plumbing here: allocation, streams,
autograd bookkeeping.  All arithmetic runs in libskp_b200.so.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import AttnLayer, call, ptr, stream_of


def _f32(t: torch.Tensor) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError("stablekeypoints_b200 kernels need CUDA tensors (there is no CPU fallback)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected float32, got {t.dtype}")
    return t.contiguous()


def _guard(t: torch.Tensor):
    return torch.cuda.device(t.device)


# --------------------------------------------------------------------------------------
# (2) readout
# --------------------------------------------------------------------------------------
def plane_argmax(maps: torch.Tensor, return_values: bool = False):
    """First flat argmax per plane (`eval.py:39-60`).  maps (N,H,W) -> int64 (N)."""
    maps = _f32(maps)
    n, h, w = maps.shape
    idx = torch.empty(n, dtype=torch.int64, device=maps.device)
    val = torch.empty(n, dtype=torch.float32, device=maps.device) if return_values else None
    with _guard(maps):
        call("skp_plane_argmax", ptr(maps), n, h, w, ptr(idx), ptr(val), stream_of(maps))
    return (idx, val) if return_values else idx


def find_k_max_pixels(maps: torch.Tensor, num: int) -> torch.Tensor:
    maps = _f32(maps)
    n, h, w = maps.shape
    out = torch.empty(num, n, 2, dtype=torch.float32, device=maps.device)
    with _guard(maps):
        call("skp_find_k_max_pixels", ptr(maps), n, h, w, num, ptr(out), stream_of(maps))
    return out


def weighted_avg_readout(maps: torch.Tensor, distance: float = 5.0, mutate: bool = False) -> torch.Tensor:
    if mutate and not (maps.is_contiguous() and maps.dtype == torch.float32):
        raise RuntimeError("in-place windowing needs a contiguous float32 tensor")
    maps = _f32(maps)
    n, h, w = maps.shape
    out = torch.empty(n, 2, dtype=torch.float32, device=maps.device)
    with _guard(maps):
        call("skp_weighted_avg_readout", ptr(maps), n, h, w, float(distance), int(mutate), ptr(out), stream_of(maps))
    return out


def plane_stats(maps: torch.Tensor, sigma: float = 2.0, epsilon: float = 1e-5, num_subjects: int = 1,
                want_kl: bool = True, want_entropy: bool = False, want_argmax: bool = False):
    """Per-plane KL(gaussian || pixel-softmax) / pixel entropy / argmax in one launch."""
    maps = _f32(maps)
    n, h, w = maps.shape
    dev = maps.device
    kl = torch.empty(n, dtype=torch.float32, device=dev) if want_kl else None
    ent = torch.empty(n, dtype=torch.float32, device=dev) if want_entropy else None
    am = torch.empty(n, dtype=torch.int64, device=dev) if want_argmax else None
    with _guard(maps):
        call("skp_plane_stats", ptr(maps), n, h, w, float(sigma), float(epsilon), int(num_subjects),
             ptr(kl), ptr(ent), ptr(am), stream_of(maps))
    return kl, ent, am


def gaussian_circles(pos: torch.Tensor, size: int, sigma: float) -> torch.Tensor:
    pos = _f32(pos)
    s, n, _ = pos.shape
    out = torch.empty(n, size, size, dtype=torch.float32, device=pos.device)
    with _guard(pos):
        call("skp_gaussian_circles", ptr(pos), s, n, size, float(sigma), ptr(out), stream_of(pos))
    return out


# --------------------------------------------------------------------------------------
# (3) selection
# --------------------------------------------------------------------------------------
def rank_ascending(score: torch.Tensor, top_k: int) -> torch.Tensor:
    score = _f32(score)
    out = torch.empty(top_k, dtype=torch.int64, device=score.device)
    with _guard(score):
        call("skp_rank_ascending", ptr(score), score.numel(), top_k, ptr(out), stream_of(score))
    return out


def rank_ascending_batched(score: torch.Tensor, top_k: int) -> torch.Tensor:
    """score (B, N) -> (B, top_k) int64: `torch.argsort(score[b])[:top_k]` for every row, one launch."""
    score = _f32(score)
    b, n = score.shape
    out = torch.empty(b, top_k, dtype=torch.int64, device=score.device)
    with _guard(score):
        call("skp_rank_ascending_batched", ptr(score), b, n, top_k, ptr(out), stream_of(score))
    return out


def furthest_point_sampling_batched(argmax_flat: torch.Tensor, h: int, w: int, candidates: torch.Tensor, top_k: int) -> torch.Tensor:
    """argmax_flat (B, N) int64, candidates (B, n_cand) unique ids per row -> (B, top_k) int64, one launch, no host sync."""
    if argmax_flat.dtype != torch.int64 or candidates.dtype != torch.int64:
        raise RuntimeError("indices must be int64")
    b, n = argmax_flat.shape
    if candidates.shape[1] < 2:
        raise NameError("furthest_pair")   # what the reference raises (ptp_utils.py:140)
    k = min(top_k, candidates.shape[1])
    out = torch.empty(b, k, dtype=torch.int64, device=argmax_flat.device)
    n_out = torch.empty(b, dtype=torch.int32, device=argmax_flat.device)
    with _guard(argmax_flat):
        call("skp_furthest_point_sampling_batched", ptr(argmax_flat.contiguous()), b, n, h, w, ptr(candidates.contiguous()),
             candidates.shape[1], k, ptr(out), ptr(n_out), stream_of(argmax_flat))
    return out


def furthest_point_sampling(argmax_flat: torch.Tensor, h: int, w: int, candidates: torch.Tensor, top_k: int,
                            assume_unique: bool = True) -> torch.Tensor:
    """`ptp_utils.py:115-159` on precomputed per-plane argmax indices.  With unique candidates
    (always true for ranked lists) exactly top_k ids come back and no host sync is needed;
    otherwise the count is read back, as the reference returns fewer ids when candidates run out."""
    if argmax_flat.dtype != torch.int64 or candidates.dtype != torch.int64:
        raise RuntimeError("indices must be int64")
    candidates = candidates.contiguous()
    if candidates.numel() < 2:
        raise NameError("furthest_pair")   # what the reference raises (ptp_utils.py:140)
    k = min(top_k, candidates.numel())
    out = torch.empty(k, dtype=torch.int64, device=argmax_flat.device)
    n_out = torch.zeros(1, dtype=torch.int32, device=argmax_flat.device)
    with _guard(argmax_flat):
        call("skp_furthest_point_sampling", ptr(argmax_flat.contiguous()), argmax_flat.numel(), h, w, ptr(candidates),
             candidates.numel(), k, ptr(out), ptr(n_out), stream_of(argmax_flat))
    if assume_unique:
        return out
    return out[: int(n_out.item())]


# --------------------------------------------------------------------------------------
# resampling
# --------------------------------------------------------------------------------------
def bilinear_upsample(x: torch.Tensor, size: int) -> torch.Tensor:
    x = _f32(x)
    lead = x.shape[:-2]
    h, w = x.shape[-2:]
    n = int(torch.tensor(lead).prod()) if len(lead) else 1
    out = torch.empty(*lead, size, size, dtype=torch.float32, device=x.device)
    with _guard(x):
        call("skp_bilinear_upsample", ptr(x), n, h, w, ptr(out), size, size, stream_of(x))
    return out


def bicubic_upsample_nhwc(x: torch.Tensor, size: int) -> torch.Tensor:
    x = _f32(x)
    b, s, s2, c = x.shape
    assert s == s2
    out = torch.empty(b, size, size, c, dtype=torch.float32, device=x.device)
    with _guard(x):
        call("skp_bicubic_upsample_nhwc", ptr(x), b, s, c, ptr(out), size, stream_of(x))
    return out


def invert_theta(theta: torch.Tensor) -> torch.Tensor:
    """Closed-form inverse of the augmented affine (`invertable_transform.py:77-84`); (..,2,3)."""
    a, b, c = theta[..., 0, 0], theta[..., 0, 1], theta[..., 0, 2]
    d, e, f = theta[..., 1, 0], theta[..., 1, 1], theta[..., 1, 2]
    det = a * e - b * d
    return torch.stack([torch.stack([e / det, -b / det, (b * f - c * e) / det], dim=-1),
                        torch.stack([-d / det, a / det, (c * d - a * f) / det], dim=-1)], dim=-2)


class _AffineWarp(torch.autograd.Function):
    """grid_sample(img, affine_grid(theta)) — `invertable_transform.py:64-68, 86-90`."""

    @staticmethod
    def forward(ctx, img, theta):
        img = _f32(img)
        theta = _f32(theta.to(img.device))
        n, c, h, w = img.shape
        out = torch.empty_like(img)
        with _guard(img):
            call("skp_affine_warp", ptr(img), ptr(theta), n, c, h, w, ptr(out), stream_of(img))
        ctx.save_for_backward(theta)
        return out

    @staticmethod
    def backward(ctx, g):
        (theta,) = ctx.saved_tensors
        g = _f32(g)
        n, c, h, w = g.shape
        d = torch.zeros_like(g)
        with _guard(g):
            call("skp_affine_warp_bwd", ptr(g), ptr(theta), n, c, h, w, ptr(d), stream_of(g))
        return d, None


def affine_warp(img: torch.Tensor, theta: torch.Tensor) -> torch.Tensor:
    return _AffineWarp.apply(img, theta)


# --------------------------------------------------------------------------------------
# (4) losses
# --------------------------------------------------------------------------------------
class _KeypointLosses(torch.autograd.Function):
    """(sharpening, equivariance) of `optimize.py:157-206` on the selected planes, fused.

    maps_a / maps_b are the FULL (T,H,W) maps; idx selects the K planes inside the kernel, so
    the (K,H,W) gathers of `optimize.py:420-424` are never materialised.  The backward writes
    compact (K,H,W) gradients and scatters them into (T,H,W) zeros (what autograd of the
    reference's `attn_map[idx]` produces)."""

    @staticmethod
    def forward(ctx, maps_a, maps_b, idx, theta_inv, sigma, num_subjects):
        maps_a = _f32(maps_a)
        t, h, w = maps_a.shape
        k = idx.numel()
        idx = idx.to(device=maps_a.device, dtype=torch.int64).contiguous()
        if maps_b is not None:
            maps_b = _f32(maps_b)
            theta_inv = _f32(theta_inv.to(maps_a.device)).reshape(2, 3)
        parts = torch.empty(k, 2, dtype=torch.float32, device=maps_a.device)
        with _guard(maps_a):
            call("skp_keypoint_losses_fwd", ptr(maps_a), ptr(maps_b), ptr(idx), k, h, w, ptr(theta_inv), float(sigma),
                 int(num_subjects), ptr(parts), stream_of(maps_a))
        ctx.save_for_backward(maps_a, maps_b, idx, theta_inv)
        ctx.cfg = (float(sigma), int(num_subjects))
        tot = parts.sum(dim=0) / float(k * h * w)
        return tot[0], tot[1]

    @staticmethod
    def backward(ctx, g_sharp, g_equiv):
        maps_a, maps_b, idx, theta_inv = ctx.saved_tensors
        sigma, ns = ctx.cfg
        t, h, w = maps_a.shape
        k = idx.numel()
        d_a = torch.empty(k, h, w, dtype=torch.float32, device=maps_a.device)
        d_b = torch.zeros(k, h, w, dtype=torch.float32, device=maps_a.device) if maps_b is not None else None
        # upstream scalars are read on the host: two floats per step
        gs, ge = float(g_sharp), float(g_equiv)
        with _guard(maps_a):
            call("skp_keypoint_losses_bwd", ptr(maps_a), ptr(maps_b), ptr(idx), k, h, w, ptr(theta_inv), sigma, ns,
                 gs, ge, ptr(d_a), ptr(d_b), stream_of(maps_a))
        ga = torch.zeros_like(maps_a).index_add_(0, idx, d_a)
        gb = torch.zeros_like(maps_b).index_add_(0, idx, d_b) if maps_b is not None else None
        return ga, gb, None, None, None, None


class _KeypointLossesBatched(torch.autograd.Function):
    """The per-image loop of `optimize.py:403-429` as one launch.  `planes` (B, K, H, W) are the selected planes of B
    images (sharpening only), or with `paired` (2B, K, H, W): rows [0, B) the images, rows [B, 2B) their transformed copies,
    theta_inv (B, 2, 3).  -> (B, 2) = per image (mean sharpening, mean equivariance).  The backward hands the (B, 2)
    upstream gradient to the kernel as a DEVICE pointer and returns one compact gradient of `planes`' shape: no host
    read, no scatter, no concatenation."""

    @staticmethod
    def forward(ctx, planes, theta_inv, paired, sigma, num_subjects):
        planes = _f32(planes)
        n, k, h, w = planes.shape
        b = n // 2 if paired else n
        dev = planes.device
        if paired:
            if n != 2 * b:
                raise RuntimeError("paired planes need an even leading dimension")
            theta_inv = _f32(theta_inv.to(dev)).reshape(b, 6)
        planes_b = planes[b:] if paired else None
        idx = torch.arange(k, device=dev, dtype=torch.int64)
        parts = torch.empty(b, k, 2, dtype=torch.float32, device=dev)
        with _guard(planes):
            call("skp_keypoint_losses_batched_fwd", ptr(planes), ptr(planes_b), ptr(idx), 0, b, k, k, h, w,
                 ptr(theta_inv) if paired else None, float(sigma), int(num_subjects), ptr(parts), stream_of(planes))
        ctx.save_for_backward(planes, idx, theta_inv if paired else None)
        ctx.cfg = (float(sigma), int(num_subjects), bool(paired))
        return parts.sum(dim=1) / float(k * h * w)

    @staticmethod
    def backward(ctx, g):
        planes, idx, theta_inv = ctx.saved_tensors
        sigma, ns, paired = ctx.cfg
        n, k, h, w = planes.shape
        b = n // 2 if paired else n
        g = _f32(g)
        d = torch.empty_like(planes)
        if paired:
            d[b:].zero_()                      # the un-warp's adjoint accumulates (atomics)
        with _guard(planes):
            call("skp_keypoint_losses_batched_bwd", ptr(planes), ptr(planes[b:]) if paired else None, ptr(idx), 0, b, k, k, h, w,
                 ptr(theta_inv) if paired else None, sigma, ns, ptr(g), ptr(d), ptr(d[b:]) if paired else None, stream_of(planes))
        return d, None, None, None, None


def keypoint_losses_batched(planes, theta_inv=None, sigma: float = 1.0, num_subjects: int = 1) -> torch.Tensor:
    """-> (B, 2): (mean sharpening loss, mean equivariance loss) of every image, one launch each way.
    theta_inv given: `planes` is (2B, K, H, W) = images then transformed copies; else (B, K, H, W), sharpening only."""
    return _KeypointLossesBatched.apply(planes, theta_inv, theta_inv is not None, sigma, num_subjects)


def keypoint_losses(maps_a, maps_b, idx, theta_inv, sigma: float, num_subjects: int = 1):
    """-> (mean sharpening loss, mean equivariance loss) for one image."""
    return _KeypointLosses.apply(maps_a, maps_b, idx, theta_inv, sigma, num_subjects)


# --------------------------------------------------------------------------------------
# (5) TTA ensemble
# --------------------------------------------------------------------------------------
def tta_accumulate(maps: torch.Tensor, theta_inv: torch.Tensor, sum_: torch.Tensor, num_: torch.Tensor) -> None:
    """maps (A,K,h,w) feature-resolution planes; sum_/num_ (K,H,W) updated in place."""
    maps = _f32(maps)
    a, k, h, w = maps.shape
    kk, H, W = sum_.shape
    assert kk == k and num_.shape == sum_.shape
    theta_inv = _f32(theta_inv.to(maps.device)).reshape(a, 6)
    with _guard(maps):
        call("skp_tta_accumulate", ptr(maps), ptr(theta_inv), a, k, h, w, ptr(sum_), ptr(num_), H, W, stream_of(maps))


def tta_finalize(sum_: torch.Tensor, num_: torch.Tensor) -> torch.Tensor:
    out = torch.empty_like(sum_)
    with _guard(sum_):
        call("skp_tta_finalize", ptr(sum_), ptr(num_), sum_.numel(), ptr(out), stream_of(sum_))
    return out


# --------------------------------------------------------------------------------------
# (1) fused attention maps
# --------------------------------------------------------------------------------------
def _layer_array(qs, ks, heads, scales, weights, dqs=None, dks=None):
    arr = (AttnLayer * len(qs))()
    for i, (q, k) in enumerate(zip(qs, ks)):
        b, n, c = q.shape
        s = int(round(n ** 0.5))
        if s * s != n:
            raise RuntimeError(f"layer {i}: q has {n} positions, not a square")
        if k.shape[-1] != c or k.shape[0] not in (1, b):
            raise RuntimeError(f"layer {i}: k shape {tuple(k.shape)} does not match q {tuple(q.shape)}")
        arr[i].q = q.data_ptr()
        arr[i].k = k.data_ptr()
        arr[i].dq = dqs[i].data_ptr() if dqs is not None else None
        arr[i].dk = dks[i].data_ptr() if dks is not None else None
        arr[i].s, arr[i].C, arr[i].heads = s, c, int(heads[i])
        arr[i].k_shared = int(k.shape[0] == 1 and b > 1)
        arr[i].scale, arr[i].weight = float(scales[i]), float(weights[i])
    return arr


def _workspace(arr, n_layers, B, T, R, prec, device, call=_lib.WS_ANY):
    # sized for the call that follows (the any-call query includes the CUDA-core backward's 2.4 GB at BASELINE shapes)
    need = _lib.load().skp_attn_map_workspace_bytes_for(arr, n_layers, B, T, R, prec, call)
    return torch.empty(max(int(need), 16), dtype=torch.uint8, device=device), int(need)


class _AttentionMaps(torch.autograd.Function):
    """maps[b,t,p] = sum_l w_l * mean_head softmax_t(scale_l q'_l k_l^T) — replaces the captured branch of the
    patched forward (`ptp_utils.py:508-538`) and `collect_maps` (`optimize.py:27-79`).  Saved for
    backward: q, k only (the reference keeps 262-655 MB of probabilities per layer alive)."""

    @staticmethod
    def forward(ctx, meta, token_idx, *tensors):
        heads, scales, weights, R, prec = meta
        n = len(heads)
        qs = [_f32(t) for t in tensors[:n]]
        ks = [_f32(t) for t in tensors[n:]]
        B, T = qs[0].shape[0], ks[0].shape[1]
        dev = qs[0].device
        stride = 0
        if token_idx is not None:
            token_idx = token_idx.to(device=dev, dtype=torch.int64).contiguous()
            if token_idx.dim() == 2:          # one row of ids per image
                if token_idx.shape[0] != B:
                    raise RuntimeError(f"token_idx has {token_idx.shape[0]} rows for {B} images")
                stride = token_idx.shape[1]
        t_out = token_idx.shape[-1] if token_idx is not None else T
        maps = torch.empty(B, t_out, R, R, dtype=torch.float32, device=dev)
        # per (image, head, pixel) softmax statistics (row max, sum exp2): 8 B per row instead of
        # the reference's T probabilities; the tcgen05 backward consumes them
        stats = torch.empty(B, sum(heads), R * R, 2, dtype=torch.float32, device=dev)
        arr = _layer_array(qs, ks, heads, scales, weights)
        with _guard(maps):
            ws, nbytes = _workspace(arr, n, B, T, R, prec, dev, _lib.WS_FWD)
            call("skp_attn_map_fwd", arr, n, B, T, R, ptr(token_idx), t_out if token_idx is not None else 0, stride,
                 ptr(maps), ptr(stats), prec, ptr(ws), nbytes, stream_of(maps))
        ctx.save_for_backward(token_idx, stats, *qs, *ks)
        ctx.meta = meta
        ctx.mark_non_differentiable(stats, ws)
        return maps, stats, ws

    @staticmethod
    def backward(ctx, g, _gstats=None, _gws=None):
        heads, scales, weights, R, prec = ctx.meta
        n = len(heads)
        token_idx, stats = ctx.saved_tensors[0], ctx.saved_tensors[1]
        qs = list(ctx.saved_tensors[2:2 + n])
        ks = list(ctx.saved_tensors[2 + n:])
        B, T = qs[0].shape[0], ks[0].shape[1]
        dev = g.device
        g = _f32(g)
        dqs = [torch.zeros_like(q) for q in qs]
        dks = [torch.zeros_like(k) for k in ks]
        chunk = max(1, min(16, min(q.shape[-1] // h for q, h in zip(qs, heads))))
        arr = _layer_array(qs, ks, heads, scales, weights, dqs, dks)
        if token_idx is not None and token_idx.dim() == 2:
            # per-image ids: the gradient is already compact, (B, n_idx, R, R)
            with _guard(g):
                ws, nbytes = _workspace(arr, n, B, T, R, prec, dev, _lib.WS_BWD_STATS if stats is not None else _lib.WS_BWD_NOSTATS)
                for i in range(0, token_idx.shape[1], chunk):
                    tok = token_idx[:, i:i + chunk].contiguous()
                    dsel = g[:, i:i + chunk].contiguous()
                    call("skp_attn_map_bwd", arr, n, B, T, R, ptr(tok), tok.shape[1], tok.shape[1], ptr(dsel), ptr(stats),
                         prec, ptr(ws), nbytes, stream_of(g))
            return (None, None, *dqs, *dks)
        # the losses touch K selected planes only (optimize.py:420-424): find them, compact them
        live = (g.abs().amax(dim=(0, 2, 3)) > 0).nonzero().flatten()
        if live.numel():
            tok = live if token_idx is None else token_idx[live]
            with _guard(g):
                ws, nbytes = _workspace(arr, n, B, T, R, prec, dev, _lib.WS_BWD_STATS if stats is not None else _lib.WS_BWD_NOSTATS)
                for i in range(0, live.numel(), chunk):
                    sel = live[i:i + chunk]
                    dsel = g.index_select(1, sel).contiguous()
                    call("skp_attn_map_bwd", arr, n, B, T, R, ptr(tok[i:i + chunk].contiguous()), sel.numel(), 0, ptr(dsel),
                         ptr(stats), prec, ptr(ws), nbytes, stream_of(g))
        return (None, None, *dqs, *dks)


def _attn_backward_compact(meta, token_idx, g, qs, ks, stats=None, fwd_ws=None):
    """dq, dk for a compact gradient g (B, n_idx, R, R) on planes token_idx (n_idx,) or (B, n_idx).
    fwd_ws: the workspace tensor of the forward of the same q / k (`attention_maps(..., return_workspace=True)`): the backward
    reads the native-resolution logits and interpolation tables the forward staged in it instead of recomputing them."""
    heads, scales, weights, R, prec = meta
    n = len(heads)
    B, T = qs[0].shape[0], ks[0].shape[1]
    g = _f32(g)
    dqs = [torch.zeros_like(q) for q in qs]
    dks = [torch.zeros_like(k) for k in ks]
    chunk = max(1, min(16, min(q.shape[-1] // h for q, h in zip(qs, heads))))
    arr = _layer_array(qs, ks, heads, scales, weights, dqs, dks)
    per_image = token_idx.dim() == 2
    with _guard(g):
        ws, nbytes = _workspace(arr, n, B, T, R, prec, g.device, _lib.WS_BWD_STATS if stats is not None else _lib.WS_BWD_NOSTATS)
        for i in range(0, token_idx.shape[-1], chunk):
            tok = token_idx[..., i:i + chunk].contiguous()
            dsel = g[:, i:i + chunk].contiguous()
            call("skp_attn_map_bwd_reuse", arr, n, B, T, R, ptr(tok), tok.shape[-1], tok.shape[-1] if per_image else 0,
                 ptr(dsel), ptr(stats), prec, ptr(ws), nbytes, ptr(fwd_ws), fwd_ws.numel() if fwd_ws is not None else 0, stream_of(g))
    return dqs, dks


class _SelectedPlanes(torch.autograd.Function):
    """maps_full[b, token_idx[b]] as a differentiable function of (q, k).

    Forward gathers K planes per image from maps that `attention_maps` already produced under
    no_grad (token selection is integer-valued, so nothing differentiates through it); backward
    is the compact attention-map backward.  This is how `optimize_embedding` avoids both a
    second forward and the dense (T,R,R) zero-padded gradients autograd would otherwise build
    for `attn_map[top_embedding_indices]` (`optimize.py:420-424`)."""

    @staticmethod
    def forward(ctx, meta, token_idx, maps_full, stats, fwd_ws, *tensors):
        n = len(meta[0])
        qs = [_f32(t) for t in tensors[:n]]
        ks = [_f32(t) for t in tensors[n:]]
        B, K = token_idx.shape
        R = meta[3]
        out = torch.gather(maps_full, 1, token_idx.view(B, K, 1, 1).expand(B, K, R, R))
        ctx.has_stats = stats is not None
        ctx.fwd_ws = fwd_ws if stats is not None else None        # a plain attribute: never part of the graph
        ctx.save_for_backward(token_idx, *([stats] if stats is not None else []), *qs, *ks)
        ctx.meta = meta
        return out

    @staticmethod
    def backward(ctx, g):
        n = len(ctx.meta[0])
        saved = list(ctx.saved_tensors)
        token_idx = saved.pop(0)
        stats = saved.pop(0) if ctx.has_stats else None
        qs, ks = saved[:n], saved[n:]
        dqs, dks = _attn_backward_compact(ctx.meta, token_idx, g, qs, ks, stats, ctx.fwd_ws)
        ctx.fwd_ws = None
        return (None, None, None, None, None, *dqs, *dks)


def _meta(heads, scales, weights, feature_res, precision, n):
    if weights is None:
        weights = [1.0 / n] * n
    return (tuple(int(h) for h in heads), tuple(float(s) for s in scales), tuple(float(w) for w in weights),
            int(feature_res), _lib.PRECISIONS[precision])


def selected_planes(maps_full: torch.Tensor, token_idx: torch.Tensor, qs, ks, heads, scales, feature_res: int,
                    weights=None, precision: str = "auto", row_stats: Optional[torch.Tensor] = None,
                    fwd_workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(B, K, R, R) = maps_full[b, token_idx[b]] with gradients flowing to qs / ks (see _SelectedPlanes).
    `row_stats` = the statistics `attention_maps(..., return_stats=True)` returned with `maps_full`; `fwd_workspace` = the
    workspace tensor of that same call (`return_workspace=True`), which lets the backward reuse the operands it holds."""
    token_idx = token_idx.to(device=maps_full.device, dtype=torch.int64).contiguous()
    return _SelectedPlanes.apply(_meta(heads, scales, weights, feature_res, precision, len(qs)), token_idx,
                                 maps_full, row_stats, fwd_workspace, *qs, *ks)


def attention_maps(qs: Sequence[torch.Tensor], ks: Sequence[torch.Tensor], heads: Sequence[int], scales: Sequence[float],
                   feature_res: int, token_idx: Optional[torch.Tensor] = None, weights: Optional[Sequence[float]] = None,
                   precision: str = "auto", return_stats: bool = False, return_workspace: bool = False):
    """qs[l] (B, s_l^2, C_l) low-resolution to_q(x); ks[l] (B or 1, T, C_l) to_k(context).
    -> (B, T', R, R) with T' = len(token_idx) or T.  Differentiable w.r.t. every q and k.
    return_stats: also return the (B, sum heads, R^2, 2) softmax statistics (for `selected_planes`).    `precision`: "auto" / "tc_bf16x3" = tcgen05 with bf16 hi + lo operands (forward fp32-grade, rtol 1e-3; the BACKWARD of
    these maps takes single-bf16 operands in its gradient products: <= 1.5e-2 of each gradient's max, measured <= 7e-3),
    "tc_bf16" = plain bf16 forward (rtol 3e-2), "simt_f32" = fp32 CUDA-core kernels in both directions (2e-3, validation).
    """
    n = len(qs)
    if weights is None:
        weights = [1.0 / n] * n
    meta = (tuple(int(h) for h in heads), tuple(float(s) for s in scales), tuple(float(w) for w in weights),
            int(feature_res), _lib.PRECISIONS[precision])
    maps, stats, ws = _AttentionMaps.apply(meta, token_idx, *qs, *ks)
    if return_workspace:
        return maps, stats, ws
    return (maps, stats) if return_stats else maps
