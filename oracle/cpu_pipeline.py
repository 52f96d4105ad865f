"""CPU restatement of the reference's END-TO-END hot path (hook -> collect_maps -> selection ->
losses -> backward; TTA extraction) on the stand-in pipeline, built from the oracle's functions.

TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Used by `bench.py`'s `cpu_baseline` leg and by
`bench.py --impl reference` (the reference itself cannot travel to the GPU box), and by tests.
It follows the reference's op sequence — including the parts the CUDA path avoids (to_q on the
upsampled hidden state, the stored (h,R^2,T) probabilities, two separate UNet passes, one image
per pass) — so that its timing is the reference algorithm's cost on host cores, kind "port".
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from . import keypoint_oracle as O


class CpuStore:
    """`AttentionStore` (`ptp_utils.py:63-83`): a list of captured (B*h, R^2, T) tensors."""

    def __init__(self, max_query_tokens: int = 32 ** 2, max_layers: int = 4):
        self.attn: List[torch.Tensor] = []
        self.max_query_tokens, self.max_layers = max_query_tokens, max_layers

    def reset(self):
        self.attn = []


def register_cpu_hook(unet, store: CpuStore, feature_res: int) -> int:
    """Restates `register_attention_control` (`ptp_utils.py:472-573`): every `CrossAttention`
    under an "up" child gets a forward that runs standard attention (`:483-506,540-541`) and, for
    the gated layers (`:508-512`), also stores the recomputed probabilities at feature_res."""

    def make(mod):
        def forward(x, context=None, mask=None):
            is_cross = context is not None
            ctx = context if is_cross else x
            q = O.split_heads(mod.to_q(x), mod.heads)
            k = O.split_heads(mod.to_k(ctx), mod.heads)
            v = O.split_heads(mod.to_v(ctx), mod.heads)
            attn = torch.softmax(torch.einsum("bid,bjd->bij", q, k) * mod.scale, dim=-1)
            out = torch.matmul(attn, v)
            if is_cross and x.shape[1] <= store.max_query_tokens and len(store.attn) < store.max_layers:
                store.attn.append(O.captured_layer_probs(x, ctx, mod.to_q.weight, mod.to_k.weight, mod.heads, feature_res))
            return mod.to_out[0](mod.reshape_batch_dim_to_heads(out))
        return forward

    n = 0
    for name, child in unet.named_children():
        if "up" in name:
            for m in child.modules():
                if m.__class__.__name__ == "CrossAttention":
                    m.forward = make(m)
                    n += 1
    assert n != 0
    return n


def cpu_maps(ldm, store: CpuStore, image, context, noise, layers=(0, 1, 2, 3), upsample_res=-1, indices=None):
    """`run_and_find_attn` (`ptp_utils.py:234-272`) for ONE image (1,3,H,W) -> (T',R',R')."""
    store.reset()
    with torch.no_grad():
        latent = ldm.vae.encode(image * 2 - 1)["latent_dist"].mean * 0.18215       # ptp_utils.py:289-304
    t = ldm.scheduler.timesteps[-1]
    noisy = ldm.scheduler.add_noise(latent, noise, t)                              # ptp_utils.py:221-223
    ldm.unet(noisy, t.repeat(1), context)                                          # ptp_utils.py:227-229
    maps = O.collect_maps(store.attn, layers=layers, upsample_res=upsample_res, indices=indices)
    store.reset()
    return maps


def cpu_token_opt_iteration(ldm, store, image, theta, context, noise_a, noise_b, *, n_cand=25, top_k=10, sigma=2.0,
                            w_sharp=100.0, w_equiv=1000.0, layers=(0, 1, 2, 3)):
    """One image-iteration of `optimize_embedding` (`optimize.py:372-445`): two UNet passes,
    selection, both losses, backward into `context` (which must require grad)."""
    maps_a = cpu_maps(ldm, store, image, context, noise_a, layers)
    transformed = O.affine_warp(image, theta[None] if theta.dim() == 2 else theta)
    maps_b = cpu_maps(ldm, store, transformed, context, noise_b, layers)
    with torch.no_grad():
        cand = O.gaussian_kl_rank(maps_a.detach(), n_cand, sigma)
        idx = O.furthest_points(maps_b.detach(), top_k, cand)
    sharp = O.sharpening_loss(maps_a[idx], sigma)
    equiv = O.equivariance_loss(maps_a[idx], maps_b[idx], theta.reshape(2, 3))
    (equiv * w_equiv + sharp * w_sharp).backward()
    return sharp.detach(), equiv.detach(), idx


@torch.no_grad()
def cpu_extract_image(ldm, store, image, context, indices, thetas: Sequence[torch.Tensor], noises, *, upscale=512,
                      layers=(0, 1, 2, 3)):
    """`run_image_with_context_augmented` + `find_max_pixel` (`eval.py:197-355, 39-60`) for one image."""
    maps, ths = [], []
    for th, nz in zip(thetas, noises):
        aug = O.affine_warp(image, th[None])
        maps.append(cpu_maps(ldm, store, aug, context, nz, layers, upsample_res=upscale, indices=indices))
        ths.append(th)
    ens = O.tta_ensemble(maps, ths)
    return ens, O.argmax_pixel(ens)
