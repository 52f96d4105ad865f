"""Attention-controller API and token selection — host mirror of the hot-path part of the
reference's `unsupervised_keypoints/ptp_utils.py` (same names, argument meaning, error
behaviour), on the CUDA kernels of libskp_b200.so.

Reference lines replaced:
  AttentionControl / AttentionStore   ptp_utils.py:32-83
  find_top_k_gaussian                 ptp_utils.py:86-112
  furthest_point_sampling             ptp_utils.py:115-159
  entropy_sort                        ptp_utils.py:165-187
  find_pred_noise / image2latent      ptp_utils.py:205-231, 289-304
  run_and_find_attn                   ptp_utils.py:234-272
  register_attention_control          ptp_utils.py:472-573
  init_random_noise                   ptp_utils.py:649-650

What changes in mechanism (never in result):
  * the captured branch of the patched forward (`:508-538`) stores a lightweight handle
    (low-resolution q = to_q(x), k = to_k(context)) instead of the (B*h, R^2, T) probability
    tensor; `collect_maps` turns the handles of the selected layers into maps with ONE fused
    kernel (bicubic q upsample, Q.K^T on tcgen05, softmax over tokens, head/layer mean).
    `handle.materialize()` still yields the reference's tensor for callers that poke at
    `controller.step_store["attn"]`.
  * registration patches once per (module, controller) instead of on every UNet forward.
"""
from __future__ import annotations

import abc
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from . import ops
from .eval import find_k_max_pixels, find_max_pixel


# --------------------------------------------------------------------------------------
# controllers (`ptp_utils.py:32-83`)
# --------------------------------------------------------------------------------------
class AttentionControl(abc.ABC):
    def step_callback(self, x_t):
        return x_t

    def between_steps(self):
        return

    @property
    def num_uncond_att_layers(self):
        return 0

    @abc.abstractmethod
    def forward(self, dict, is_cross: bool, place_in_unet: str):
        raise NotImplementedError

    def __call__(self, dict, is_cross: bool, place_in_unet: str):
        dict = self.forward(dict, is_cross, place_in_unet)
        return dict["attn"]

    def reset(self):
        self.cur_step = 0
        self.cur_att_layer = 0

    def __init__(self):
        self.cur_step = 0
        self.num_att_layers = -1
        self.cur_att_layer = 0


class CapturedAttention:
    """What the patched forward hands to the controller for one captured layer.

    Stands in for the reference's `softmax(q' k^T * scale)` tensor of shape (B*h, R^2, T)
    (`ptp_utils.py:531-536`); `shape` reports that shape, `materialize()` computes it."""

    __slots__ = ("q", "k", "heads", "scale", "side", "feature_res", "k_shared")

    def __init__(self, q, k, heads, scale, side, feature_res, k_shared):
        self.q, self.k = q, k                 # (B, side^2, C) ; (B or 1, T, C)
        self.heads, self.scale = int(heads), float(scale)
        self.side, self.feature_res = int(side), int(feature_res)
        self.k_shared = bool(k_shared)

    @property
    def shape(self):
        return (self.q.shape[0] * self.heads, self.feature_res ** 2, self.k.shape[1])

    def materialize(self) -> torch.Tensor:
        """The reference's stored tensor, for inspection/debugging (not on the hot path)."""
        b, n, c = self.q.shape
        h, d = self.heads, c // self.heads
        R = self.feature_res
        q_up = ops.bicubic_upsample_nhwc(self.q.detach().reshape(b, self.side, self.side, c), R).reshape(b, R * R, c)
        k = self.k.detach().expand(b, -1, -1)
        qh = q_up.reshape(b, R * R, h, d).permute(0, 2, 1, 3).reshape(b * h, R * R, d)
        kh = k.reshape(b, -1, h, d).permute(0, 2, 1, 3).reshape(b * h, -1, d)
        return torch.softmax(torch.einsum("bid,bjd->bij", qh, kh) * self.scale, dim=-1)


class AttentionStore(AttentionControl):
    """`ptp_utils.py:63-83`.  `step_store["attn"]` is a list, one entry per captured layer.

    Two knobs generalise the reference's hard-coded capture gate (`ptp_utils.py:508-512`):
    `max_query_tokens` (32**2) and `max_captured_layers` (4); BASELINE config 4 ("64x64 layers
    included") sets them to 64**2 and 9.  `stop_after_capture` lets the UNet stand-in return as
    soon as the last needed layer has been captured (the predicted noise is discarded by
    `run_and_find_attn`, `ptp_utils.py:246`)."""

    @staticmethod
    def get_empty_store():
        return {"attn": []}

    def forward(self, dict, is_cross: bool, place_in_unet: str):
        self.step_store["attn"].append(dict["attn"])
        return dict

    def reset(self):
        super(AttentionStore, self).reset()
        self.step_store = self.get_empty_store()

    def __init__(self, max_query_tokens: int = 32 ** 2, max_captured_layers: int = 4):
        super(AttentionStore, self).__init__()
        self.step_store = self.get_empty_store()
        self.max_query_tokens = max_query_tokens
        self.max_captured_layers = max_captured_layers
        self.stop_after_capture = False


# --------------------------------------------------------------------------------------
# registration (`ptp_utils.py:472-573`)
# --------------------------------------------------------------------------------------
class StopForward(Exception):
    """Raised inside the patched forward once the controller holds every layer it will keep."""


def register_attention_control(model, controller, feature_upsample_res=256):
    def ca_forward(self, place_in_unet):
        to_out = self.to_out
        if type(to_out) is torch.nn.modules.container.ModuleList:
            to_out = self.to_out[0]

        def forward(x, context=None, mask=None):
            batch_size, sequence_length, dim = x.shape
            h = self.heads
            is_cross = context is not None
            store = getattr(controller, "step_store", None)
            capture = (is_cross and store is not None
                       and sequence_length <= getattr(controller, "max_query_tokens", 32 ** 2)
                       and len(store["attn"]) < getattr(controller, "max_captured_layers", 4))
            q = self.to_q(x)
            # one context for the whole batch (`context.expand(B, -1, -1)`, what `ptp_utils.py:229`'s repeat amounts
            # to): detected from the tensor itself, per call, so a genuinely per-sample (B, T, C) context (CFG,
            # SDXL-style conditioning) through the same controller is never collapsed to sample 0's
            shared = is_cross and batch_size > 1 and context.shape[0] == batch_size and context.stride(0) == 0
            ctx = (context[:1] if shared else context) if is_cross else x
            k = self.to_k(ctx)
            if capture:
                side = int(sequence_length ** 0.5)
                controller({"attn": CapturedAttention(q, k, h, self.scale, side, feature_upsample_res, shared)},
                           is_cross, place_in_unet)
                if getattr(controller, "stop_after_capture", False) and \
                        len(store["attn"]) >= getattr(controller, "max_captured_layers", 4):
                    raise StopForward()
            v = self.to_v(ctx)
            d = q.shape[-1] // h
            qh = q.reshape(batch_size, -1, h, d).transpose(1, 2)
            kh = k.reshape(k.shape[0], -1, h, d).transpose(1, 2).expand(batch_size, -1, -1, -1)
            vh = v.reshape(v.shape[0], -1, h, d).transpose(1, 2).expand(batch_size, -1, -1, -1)
            am = None
            if mask is not None:
                am = mask.reshape(batch_size, 1, 1, -1)
            out = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=am, scale=self.scale)
            out = out.transpose(1, 2).reshape(batch_size, -1, h * d)
            return to_out(out)

        return forward

    class DummyController:
        def __call__(self, *args):
            return args[0]

        def __init__(self):
            self.num_att_layers = 0

    if controller is None:
        controller = DummyController()

    def register_recr(net_, count, place_in_unet):
        if net_.__class__.__name__ == "CrossAttention":
            net_.forward = ca_forward(net_, place_in_unet)
            return count + 1
        elif hasattr(net_, "children"):
            for net__ in net_.children():
                count = register_recr(net__, count, place_in_unet)
        return count

    cross_att_count = 0
    for net in model.named_children():
        if "up" in net[0]:
            cross_att_count += register_recr(net[1], 0, "up")

    controller.num_att_layers = cross_att_count
    assert cross_att_count != 0, \
        "No cross attention layers found in the model. Please check to make sure you're using diffusers==0.8.0."


# --------------------------------------------------------------------------------------
# UNet pass (`ptp_utils.py:205-272, 289-304`)
# --------------------------------------------------------------------------------------
def init_random_noise(device, num_words=77, dim=768):
    """`ptp_utils.py:649-650` (the reference hard-codes 768; SD-2.1 needs 1024, SDXL 2048)."""
    return torch.randn(1, num_words, dim).to(device)


def image2latent(model, image, device):
    """`ptp_utils.py:289-304`: (B,H,W,3) in [0,1] -> VAE mean * 0.18215.  A 4-D tensor that is
    already latent-shaped (4 channels) passes through, as in the reference."""
    with torch.no_grad():
        if torch.is_tensor(image) and image.dim() == 4 and image.shape[1] == 4:
            return image
        if not torch.is_tensor(image):
            image = torch.as_tensor(image).permute(0, 3, 1, 2)     # numpy NHWC, the reference's format
        image = image.to(device=device, dtype=torch.float32) * 2 - 1
        vae = model.vae.module if hasattr(model.vae, "module") else model.vae
        latents = vae.encode(image)["latent_dist"].mean
        return latents * 0.18215


def find_pred_noise(ldm, image, context, noise_level=-1, device="cuda", noise=None, generator=None):
    """`ptp_utils.py:205-231`.  `image` (B,3,H,W) in [0,1] stays on the GPU (the reference goes
    through numpy).  `noise` pins `randn_like(latent)` for parity tests; `generator` (a torch.Generator
    on the latent's device) draws it from a private stream (per-rank / per-image seeding)."""
    latent = image2latent(ldm, image, device)
    if noise is None:
        noise = (torch.randn_like(latent) if generator is None
                 else torch.randn(latent.shape, device=latent.device, dtype=latent.dtype, generator=generator))
    t = ldm.scheduler.timesteps[noise_level]
    noisy_image = ldm.scheduler.add_noise(latent, noise.to(latent.device), t)
    b = noisy_image.shape[0]
    try:
        pred = ldm.unet(noisy_image, t.repeat(b).to(latent.device), context.expand(b, -1, -1))["sample"]
    except StopForward:      # every layer the controller keeps has been captured; the rest is discarded anyway
        pred = None
    return noise, pred


def _controller_list(controllers) -> List[AttentionStore]:
    if isinstance(controllers, dict):
        return list(controllers.values())
    return list(controllers)


def run_and_find_attn_batched(ldm, image, context, noise_level=-1, device="cuda",
                              from_where=["down_cross", "mid_cross", "up_cross"], layers=[0, 1, 2, 3, 4, 5],
                              upsample_res=32, indices=None, controllers=None, noise=None, generator=None) -> torch.Tensor:
    """One UNet pass over a batch of images on this device -> (B, T', R', R') maps (one fused
    kernel for all images).  `run_and_find_attn` is the reference-shaped view of this."""
    from .optimize import collect_maps_batched
    ctrl = _controller_list(controllers)[0]
    ctrl.reset()
    find_pred_noise(ldm, image, context, noise_level=noise_level, device=device, noise=noise, generator=generator)
    maps = collect_maps_batched(ctrl, from_where=from_where, upsample_res=upsample_res, layers=layers, indices=indices)
    ctrl.reset()
    return maps


def run_and_find_attn(ldm, image, context, noise_level=-1, device="cuda",
                      from_where=["down_cross", "mid_cross", "up_cross"], layers=[0, 1, 2, 3, 4, 5],
                      upsample_res=32, indices=None, controllers=None, noise=None):
    """`ptp_utils.py:234-272` -> list of (T',R',R') maps, one per image of the batch.

    The reference returns one map per DataParallel replica, each replica holding exactly one
    image (`optimize.py:356`); here the B images of the batch live on ONE GPU
    (one process per GPU), so the list has B entries — the same list the reference's caller
    iterates over (`optimize.py:403`)."""
    maps = run_and_find_attn_batched(ldm, image, context, noise_level, device, from_where, layers, upsample_res,
                                     indices, controllers, noise)
    return list(maps.unbind(0))


# --------------------------------------------------------------------------------------
# token selection (`ptp_utils.py:86-187`)
# --------------------------------------------------------------------------------------
def find_top_k_gaussian(attention_maps, top_k, sigma=3, epsilon=1e-5, num_subjects=1):
    """Rank tokens by KL(gaussian at argmax || softmax over pixels), ascending (`:86-112`).
    One pass over the maps (`skp_plane_stats`) + one ranking kernel; int64 on maps.device."""
    kl, _, _ = ops.plane_stats(attention_maps, sigma=float(sigma), epsilon=float(epsilon),
                               num_subjects=int(num_subjects), want_kl=True)
    return ops.rank_ascending(kl, min(int(top_k), kl.numel()))


def entropy_sort(attention_maps, top_k, min_dist=0.05):
    """Rank tokens by the entropy of their pixel-softmax, ascending (`:165-187`)."""
    _, ent, _ = ops.plane_stats(attention_maps, want_kl=False, want_entropy=True)
    return ops.rank_ascending(ent, min(int(top_k), ent.numel()))


def furthest_point_sampling(attention_maps, top_k, top_initial_candidates):
    """`ptp_utils.py:115-159`: greedy furthest-point sampling over the candidates' argmax
    pixels, strict `>` so the first candidate wins ties.  One single-block kernel instead of
    ~500 host-synchronised steps; raises NameError on < 2 candidates like the reference."""
    n, h, w = attention_maps.shape
    am = ops.plane_argmax(attention_maps)
    cand = torch.as_tensor(top_initial_candidates).to(device=attention_maps.device, dtype=torch.int64)
    return ops.furthest_point_sampling(am, h, w, cand, int(top_k))
