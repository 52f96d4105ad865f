"""Pipeline set-up and Gaussian targets — host mirror of the hot-path part of the reference's
`unsupervised_keypoints/optimize_token.py`.

Reference lines replaced:
  load_ldm          optimize_token.py:24-79
  gaussian_circle   optimize_token.py:204-224
  gaussian_circles  optimize_token.py:226-242
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn as nn

from . import ops


def load_ldm(device, type="CompVis/stable-diffusion-v1-4", feature_upsample_res=256, images_per_pass: int = 1,
             max_query_tokens: int = 32 ** 2, max_captured_layers: int = 4, stop_after_capture: bool = True,
             seed: int = 0, ldm=None):
    """`optimize_token.py:24-79` -> (ldm, controllers, num_gpus).

    * `diffusers` / checkpoints are unavailable offline, so `type` selects a random-init stand-in
      of the published architecture (`sd_standin.unet_config`: "sd15", "sd21-base", "sdxl", or the
      hub names the reference uses); pass `ldm=` to wrap an existing pipeline object instead.
    * One process drives ONE GPU: `controllers` has a single entry keyed by this device, the
      UNet/VAE get a thin `.module` wrapper (the reference reads `ldm.vae.module`,
      `ptp_utils.py:299-300`) instead of `nn.DataParallel`, and the attention hook is installed
      once here instead of from a forward_pre_hook on every forward (`:59-68`).
    * `num_gpus` keeps the reference's meaning for its callers — images per UNet pass — and is
      `images_per_pass` (the reference: one image per DataParallel replica)."""
    from . import ptp_utils
    from .sd_standin import PipelineStandIn, _ModuleWrap, unet_config
    if str(device) == "cpu":
        raise RuntimeError("stablekeypoints_b200 has no CPU path: load_ldm needs a CUDA device")
    dev = torch.device(device)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    if ldm is None:
        ldm = PipelineStandIn(unet_config(type), seed=seed).to(dev)
    unet = ldm.unet.module if hasattr(ldm.unet, "module") else ldm.unet
    vae = ldm.vae.module if hasattr(ldm.vae, "module") else ldm.vae
    ldm.unet = _ModuleWrap(unet, [dev.index])
    ldm.vae = _ModuleWrap(vae, [dev.index])
    controller = ptp_utils.AttentionStore(max_query_tokens=max_query_tokens, max_captured_layers=max_captured_layers)
    controller.stop_after_capture = stop_after_capture
    controllers: Dict[torch.device, ptp_utils.AttentionStore] = {dev: controller}
    ptp_utils.register_attention_control(unet, controller, feature_upsample_res=feature_upsample_res)
    for m in (ldm.vae, ldm.text_encoder, ldm.unet):
        for p in m.parameters():
            p.requires_grad = False
    return ldm, controllers, int(images_per_pass)


def gaussian_circle(pos, size=64, sigma=16, device="cuda"):
    """`optimize_token.py:204-224`: pos (N,2) in [0,1] (row, col) -> (N,size,size), peak 1."""
    return ops.gaussian_circles(pos.to(device=device, dtype=torch.float32)[None].contiguous(), int(size), float(sigma))


def gaussian_circles(pos, size=64, sigma=16, device="cuda"):
    """`optimize_token.py:226-242`: pos (S,N,2); mean over the S subjects -> (N,size,size)."""
    return ops.gaussian_circles(pos.to(device=device, dtype=torch.float32).contiguous(), int(size), float(sigma))


def to_channels_last(ldm):
    """Stores the UNet's / VAE's convolution weights in NHWC (`torch.channels_last`); activations follow from the first
    convolution on.  cuDNN's tensor-core convolutions compute in NHWC: with the default NCHW tensors every convolution
    of the stock trunk is bracketed by `nchwToNhwc` / `nhwcToNchw` copies (7.2 % of the device time of an optimisation
    step, profiles/r1e_launches.txt).  A memory-layout setting of stock PyTorch: no arithmetic changes (SURVEY.md §8(f)
    rank 1, caller-side UNet economy; the reference itself runs NCHW, `optimize_token.py:36-44`)."""
    for m in (ldm.unet, ldm.vae):
        m.to(memory_format=torch.channels_last)
    return ldm
