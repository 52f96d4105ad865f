"""Keypoint readout and test-time-augmentation ensemble — host mirror of the reference's
`unsupervised_keypoints/eval.py` for the hot path (same names, argument meaning and return
layouts), running on the CUDA kernels of libskp_b200.so.

Reference lines replaced:
  find_max_pixel                    eval.py:39-60
  find_k_max_pixels / mask_radius   eval.py:62-111
  pixel_from_weighted_avg           eval.py:113-155
  run_image_with_context_augmented  eval.py:197-355
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .invertable_transform import RandomAffineWithInverse


def find_max_pixel(map: torch.Tensor) -> torch.Tensor:
    """map (N,h,w) -> float (N,2) = (row+0.5, col+0.5) of the first maximum (`eval.py:39-60`)."""
    n, h, w = map.shape
    idx = ops.plane_argmax(map)
    return torch.stack([idx // w, idx % w], dim=-1).to(torch.float32) + 0.5


def mask_radius(map: torch.Tensor, max_coords: torch.Tensor, radius: float) -> torch.Tensor:
    """`eval.py:83-111`: zero every pixel with (col-c)^2 + (row-r)^2 <= radius^2 (integer pixel
    coordinates against the +0.5 centre).  Elementwise mask on device tensors; only the
    argmax search of the callers is a kernel, so this stays a two-line torch expression."""
    n, h, w = map.shape
    ys = torch.arange(h, device=map.device).view(1, h, 1)
    xs = torch.arange(w, device=map.device).view(1, 1, w)
    d2 = (xs - max_coords[:, 1].view(n, 1, 1)) ** 2 + (ys - max_coords[:, 0].view(n, 1, 1)) ** 2
    return map * (d2 > radius ** 2).float()


def find_k_max_pixels(map: torch.Tensor, num: int = 3) -> torch.Tensor:
    """`eval.py:62-81`: `num` successive maxima with a 0.05*h disc suppressed in between.
    -> (num, N, 2).  One kernel, no masked copies of the maps."""
    return ops.find_k_max_pixels(map, num)


def pixel_from_weighted_avg(heatmaps: torch.Tensor, distance: float = 5) -> torch.Tensor:
    """`eval.py:113-155` windowed soft-argmax -> (N,2) (row, col).  Like the reference this
    zeroes `heatmaps` outside the window IN PLACE (`eval.py:137`) when it can."""
    inplace = heatmaps.is_contiguous() and heatmaps.dtype == torch.float32
    return ops.weighted_avg_readout(heatmaps, float(distance), mutate=inplace)


@torch.no_grad()
def run_image_with_context_augmented(
    ldm,
    image,
    context,
    indices,
    device="cuda",
    from_where=["down_cross", "mid_cross", "up_cross"],
    layers=[0, 1, 2, 3, 4, 5],
    augmentation_iterations=20,
    noise_level=-1,
    augment_degrees=30,
    augment_scale=(0.9, 1.1),
    augment_translate=(0.1, 0.1),
    visualize=False,
    controllers=None,
    num_gpus=1,
    save_folder="outputs",
    upscale_size=512,
    thetas: Optional[torch.Tensor] = None,
    noise: Optional[torch.Tensor] = None,
):
    """TTA ensemble of `eval.py:197-355`: mean over augmentations of the un-warped maps of the
    selected tokens, normalised by coverage, NaN -> 0.  -> (K, upscale_size, upscale_size).

    Differences in mechanism, not in result: the image never leaves the GPU (the reference
    round-trips through numpy, `eval.py:241-245`); each forward carries `num_gpus` augmented
    copies (the reference's DataParallel scatter == our per-device batch); the (K,512,512)
    bilinear upsample, the inverse warp and the accumulation are one kernel
    (`skp_tta_accumulate`) reading the (K,R,R) feature-resolution planes.
    `thetas` ((iters*num_gpus,2,3)) / `noise` pin the otherwise unseeded randomness for parity tests.
    """
    from . import ptp_utils

    if visualize:
        raise NotImplementedError("visualisation (matplotlib) is outside the hot path")
    if not torch.is_tensor(image):
        image = torch.as_tensor(image)
    image = image.to(device=device, dtype=torch.float32)
    if image.dim() == 3 and image.shape[-1] == 3 and image.shape[0] != 3:
        image = image.permute(2, 0, 1)           # numpy HWC (what the reference passes) -> CHW
    k = len(indices)
    indices = torch.as_tensor(indices).to(device=device, dtype=torch.int64)
    sum_samples = torch.zeros(k, upscale_size, upscale_size, device=device)
    num_samples = torch.zeros(k, upscale_size, upscale_size, device=device)
    transform = RandomAffineWithInverse(degrees=augment_degrees, scale=augment_scale, translate=augment_translate)
    n_it = augmentation_iterations // num_gpus
    for i in range(n_it):
        batch = image[None].expand(num_gpus, -1, -1, -1)
        th = None if thetas is None else thetas[i * num_gpus:(i + 1) * num_gpus]
        augmented = transform(batch, theta=th)
        nz = None if noise is None else noise[i * num_gpus:(i + 1) * num_gpus]
        maps = ptp_utils.run_and_find_attn_batched(
            ldm, augmented, context, layers=layers, noise_level=noise_level, from_where=from_where,
            upsample_res=-1, device=device, controllers=controllers, indices=indices, noise=nz)   # (G,K,R,R)
        ops.tta_accumulate(maps, transform.inverse_theta(), sum_samples, num_samples)
    return ops.tta_finalize(sum_samples, num_samples)
