"""`torch.ops.skp_b200.*`: the hot path's entry points as registered PyTorch operators (the operator-library
face of the drop-in boundary, next to the C ABI of `include/skp_b200.h` that they call through `_lib`).

Registered with `torch.library` for the CUDA dispatch key only: there is no CPU kernel behind any of them, a CPU tensor fails
with PyTorch's own "not implemented for CPU" error.  Each operator is the forward of the corresponding function in `ops.py`
(same argument meaning; reference lines in `INTEGRATION.md` §3); gradients flow through the `torch.autograd.Function`s of
`ops.py`, which these operators do not replace.  Meta ("fake tensor") implementations give shapes and dtypes, so the operators
can sit inside graphs traced by `torch.export` / `make_fx`.

    import stablekeypoints_b200.torch_ops                # registers the library once
    maps = torch.ops.skp_b200.attention_maps(qs, ks, heads, scales, 128, None)
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import ops as _ops

_lib_def = torch.library.Library("skp_b200", "DEF")
_lib_cuda = torch.library.Library("skp_b200", "IMPL", "CUDA")
_lib_meta = torch.library.Library("skp_b200", "IMPL", "Meta")

SCHEMAS = {
    "attention_maps": "(Tensor[] qs, Tensor[] ks, int[] heads, float[] scales, int feature_res, Tensor? token_idx) -> Tensor",
    "plane_argmax": "(Tensor maps) -> Tensor",
    "find_k_max_pixels": "(Tensor maps, int num) -> Tensor",
    "weighted_avg_readout": "(Tensor maps, float distance) -> Tensor",
    "plane_stats": "(Tensor maps, float sigma, float epsilon, int num_subjects) -> (Tensor, Tensor, Tensor)",
    "rank_ascending": "(Tensor score, int top_k) -> Tensor",
    "furthest_point_sampling": "(Tensor argmax_flat, int h, int w, Tensor candidates, int top_k) -> Tensor",
    "bilinear_upsample": "(Tensor x, int size) -> Tensor",
    "bicubic_upsample_nhwc": "(Tensor x, int size) -> Tensor",
    "affine_warp": "(Tensor img, Tensor theta) -> Tensor",
}
for _name, _schema in SCHEMAS.items():
    _lib_def.define(_name + _schema)


def _attention_maps(qs: List[torch.Tensor], ks: List[torch.Tensor], heads: List[int], scales: List[float], feature_res: int,
                    token_idx: Optional[torch.Tensor]) -> torch.Tensor:
    with torch.no_grad():
        return _ops.attention_maps([q.detach() for q in qs], [k.detach() for k in ks], heads, scales, feature_res, token_idx=token_idx)


def _attention_maps_meta(qs, ks, heads, scales, feature_res, token_idx):
    n = ks[0].shape[1] if token_idx is None else token_idx.shape[-1]
    return qs[0].new_empty((qs[0].shape[0], n, feature_res, feature_res), dtype=torch.float32)


def _plane_stats(maps, sigma, epsilon, num_subjects):
    # (N,H,W) planes -> Gaussian-KL score (`find_top_k_gaussian`), pixel entropy (`entropy_sort`), flat argmax: one launch
    return _ops.plane_stats(maps, sigma=sigma, epsilon=epsilon, num_subjects=num_subjects, want_kl=True, want_entropy=True,
                            want_argmax=True)


def _plane_stats_meta(maps, sigma, epsilon, num_subjects):
    n = maps.shape[0]
    return (maps.new_empty((n,), dtype=torch.float32), maps.new_empty((n,), dtype=torch.float32),
            maps.new_empty((n,), dtype=torch.int64))


_IMPLS = {
    "attention_maps": (_attention_maps, _attention_maps_meta),
    "plane_argmax": (lambda maps: _ops.plane_argmax(maps),
                     lambda maps: maps.new_empty((maps.shape[0],), dtype=torch.int64)),
    "find_k_max_pixels": (lambda maps, num: _ops.find_k_max_pixels(maps, num),
                          lambda maps, num: maps.new_empty((num, maps.shape[0], 2), dtype=torch.float32)),
    "weighted_avg_readout": (lambda maps, distance: _ops.weighted_avg_readout(maps, distance),
                             lambda maps, distance: maps.new_empty((maps.shape[0], 2), dtype=torch.float32)),
    "plane_stats": (_plane_stats, _plane_stats_meta),
    "rank_ascending": (lambda score, top_k: _ops.rank_ascending(score, top_k),
                       lambda score, top_k: score.new_empty((top_k,), dtype=torch.int64)),
    "furthest_point_sampling": (lambda a, h, w, c, k: _ops.furthest_point_sampling(a, h, w, c, k),
                                lambda a, h, w, c, k: a.new_empty((min(k, c.numel()),), dtype=torch.int64)),
    "bilinear_upsample": (lambda x, size: _ops.bilinear_upsample(x, size),
                          lambda x, size: x.new_empty(tuple(x.shape[:-2]) + (size, size), dtype=torch.float32)),
    "bicubic_upsample_nhwc": (lambda x, size: _ops.bicubic_upsample_nhwc(x, size),
                              lambda x, size: x.new_empty((x.shape[0], size, size, x.shape[3]), dtype=torch.float32)),
    "affine_warp": (lambda img, theta: _ops.affine_warp(img.detach(), theta.detach()),
                    lambda img, theta: torch.empty_like(img)),
}
for _name, (_cuda, _meta) in _IMPLS.items():
    _lib_cuda.impl(_name, _cuda)
    _lib_meta.impl(_name, _meta)
