"""Keypoint readout and test-time-augmentation ensemble — host mirror of the reference's
`unsupervised_keypoints/eval.py` for the hot path (same names, argument meaning and return
layouts), running on the CUDA kernels of libskp_b200.so.

Reference lines replaced:
  find_max_pixel                    eval.py:39-60
  find_k_max_pixels / mask_radius   eval.py:62-111
  pixel_from_weighted_avg           eval.py:113-155
  run_image_with_context_augmented  eval.py:197-355
  swap_points / evaluate            eval.py:360-539   (metrics: SURVEY.md §8(f) rank 2)
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
def find_corresponding_points(maps: torch.Tensor, num_points: int = 10):
    """`eval.py:159-195` (unused by `main`): rank tokens by their pixel-softmax entropy summed over the
    images, keep the `num_points` lowest, return their argmax pixels per image and the token ids.
    maps (num_images, num_tokens, h, w) -> ((num_images, num_points, 2), (num_points,) int64).
    Entropies and argmaxes come from `skp_plane_stats` / `skp_plane_argmax`, the ranking from `skp_rank_ascending`."""
    ni, nt, h, w = maps.shape
    _, ent, _ = ops.plane_stats(maps.reshape(ni * nt, h, w), want_kl=False, want_entropy=True)
    order = ops.rank_ascending(ent.reshape(ni, nt).sum(dim=0), int(num_points))
    chosen = maps[:, order].reshape(ni * order.numel(), h, w).contiguous()
    return find_max_pixel(chosen).reshape(ni, order.numel(), 2), order


def _image_chw(image, device):
    if not torch.is_tensor(image):
        image = torch.as_tensor(image)
    image = image.to(device=device, dtype=torch.float32, non_blocking=True)
    if image.dim() == 3 and image.shape[-1] == 3 and image.shape[0] != 3:
        image = image.permute(2, 0, 1)           # numpy HWC (what the reference passes) -> CHW
    return image


def run_images_with_context_augmented(
    ldm,
    images,
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
    controllers=None,
    num_gpus=1,
    upscale_size=512,
    thetas: Optional[torch.Tensor] = None,
    noise: Optional[torch.Tensor] = None,
    seeds=None,
):
    """`run_image_with_context_augmented` (`eval.py:197-355`) for n images at once -> (n, K, upscale_size, upscale_size).

    Every UNet pass carries `num_gpus` augmented copies of EACH of the n images (n * num_gpus images per pass, one fused
    attention-map launch for all of them); images are independent, so the result per image is what the single-image
    call returns.  `thetas` (n, iters*num_gpus, 2, 3) / `noise` (n, iters*num_gpus, 4, h, w) pin the randomness;
    `seeds` (n ints) instead draws image j's transforms and noise from private generators seeded seeds[j], which makes
    the result independent of how images are grouped into passes or sharded over ranks."""
    from . import ptp_utils

    images = [_image_chw(im, device) for im in images]
    n = len(images)
    k = len(indices)
    indices = torch.as_tensor(indices).to(device=device, dtype=torch.int64)
    sum_samples = torch.zeros(n, k, upscale_size, upscale_size, device=device)
    num_samples = torch.zeros(n, k, upscale_size, upscale_size, device=device)
    transform = RandomAffineWithInverse(degrees=augment_degrees, scale=augment_scale, translate=augment_translate)
    g_cpu = g_dev = [None] * n
    if seeds is not None:
        g_cpu = [torch.Generator().manual_seed(int(s)) for s in seeds]
        g_dev = [torch.Generator(device=device).manual_seed(int(s)) for s in seeds]
    stacked = torch.stack(images)
    n_it = augmentation_iterations // num_gpus
    for i in range(n_it):
        sl = slice(i * num_gpus, (i + 1) * num_gpus)
        batch = stacked[:, None].expand(-1, num_gpus, -1, -1, -1).reshape(n * num_gpus, *stacked.shape[1:])
        if thetas is not None:
            th = thetas[:, sl].reshape(n * num_gpus, 2, 3)
        else:
            th = torch.cat([transform.sample_theta(num_gpus, g_cpu[j]) for j in range(n)])
        augmented = transform(batch, theta=th)
        if noise is not None:
            nz = noise[:, sl].reshape(n * num_gpus, *noise.shape[2:])
        elif seeds is not None:
            lat = (num_gpus, 4, stacked.shape[-2] // 8, stacked.shape[-1] // 8)
            nz = torch.cat([torch.randn(lat, device=device, generator=g_dev[j]) for j in range(n)])
        else:
            nz = None
        maps = ptp_utils.run_and_find_attn_batched(
            ldm, augmented, context, layers=layers, noise_level=noise_level, from_where=from_where,
            upsample_res=-1, device=device, controllers=controllers, indices=indices, noise=nz)   # (n*G,K,R,R)
        th_inv = transform.inverse_theta()
        for j in range(n):
            ops.tta_accumulate(maps[j * num_gpus:(j + 1) * num_gpus], th_inv[j * num_gpus:(j + 1) * num_gpus],
                               sum_samples[j], num_samples[j])
    return ops.tta_finalize(sum_samples, num_samples)


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
    seed: Optional[int] = None,
):
    """TTA ensemble of `eval.py:197-355`: mean over augmentations of the un-warped maps of the
    selected tokens, normalised by coverage, NaN -> 0.  -> (K, upscale_size, upscale_size).

    Differences in mechanism, not in result: the image never leaves the GPU (the reference
    round-trips through numpy, `eval.py:241-245`); each forward carries `num_gpus` augmented
    copies (the reference's DataParallel scatter == our per-device batch); the (K,512,512)
    bilinear upsample, the inverse warp and the accumulation are one kernel
    (`skp_tta_accumulate`) reading the (K,R,R) feature-resolution planes.
    `thetas` ((iters*num_gpus,2,3)) / `noise` pin the otherwise unseeded randomness for parity tests;
    `seed` draws both from private generators.
    """
    if visualize:
        raise NotImplementedError("visualisation (matplotlib) is outside the hot path")
    return run_images_with_context_augmented(
        ldm, [image], context, indices, device=device, from_where=from_where, layers=layers,
        augmentation_iterations=augmentation_iterations, noise_level=noise_level, augment_degrees=augment_degrees,
        augment_scale=augment_scale, augment_translate=augment_translate, controllers=controllers, num_gpus=num_gpus,
        upscale_size=upscale_size, thetas=None if thetas is None else thetas[None], noise=None if noise is None else noise[None],
        seeds=None if seed is None else [seed])[0]


# --------------------------------------------------------------------------------------
# evaluation (`eval.py:360-539`)
# --------------------------------------------------------------------------------------
_H36M_PAIRS = ((1, 6), (2, 7), (3, 8), (4, 9), (5, 10), (17, 25), (18, 26), (19, 27), (20, 28), (21, 28), (22, 30), (23, 31))


def swap_points(points: torch.Tensor) -> torch.Tensor:
    """Left/right relabelling of Human3.6M keypoints, points (B, N, D) (`eval.py:360-371`; the table,
    including its (21, 28) entry, is the reference's)."""
    perm = list(range(points.shape[1]))
    for a, b in _H36M_PAIRS:
        perm[a] = b
        perm[b] = a
    return points[:, perm, :]


def keypoint_error(estimated_kpts: torch.Tensor, gt_kpts: torch.Tensor, evaluation_method: str = "inter_eye_distance",
                   visible: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-image error of `evaluate` (`eval.py:469-510`) for (n, 2) regressed / annotated keypoints in [0, 1]."""
    if evaluation_method in ("mean_average_error", "pck"):
        estimated_kpts, gt_kpts = estimated_kpts * 256, gt_kpts * 256
    l2 = (estimated_kpts - gt_kpts).norm(dim=-1)
    if evaluation_method == "inter_eye_distance":
        return (l2 / torch.sqrt(((gt_kpts[0] - gt_kpts[1]) ** 2).sum(dim=-1))).mean()
    if evaluation_method in ("visible", "mean_average_error"):
        vis = torch.ones_like(l2) if visible is None else visible.to(l2)
        err = (l2 * vis).sum()
        return err / vis.sum() if evaluation_method == "visible" else err
    if evaluation_method == "pck":
        return (l2 < 6).float().mean()
    if evaluation_method == "orientation_invariant":
        swapped = (swap_points(estimated_kpts[None])[0] - gt_kpts).norm(dim=-1).mean()
        return torch.minimum(l2.mean(), swapped) * 128
    raise NotImplementedError(evaluation_method)


@torch.no_grad()
def extract_keypoints(ldm, samples, context, indices, *, device, from_where, layers, noise_level, augmentation_iterations,
                      augment_degrees, augment_scale, augment_translate, controllers, num_gpus, max_loc_strategy="argmax",
                      seeds=None, upscale_size=512):
    """Keypoints in [0, 1] of a group of dataset samples: TTA maps of all of them in shared UNet passes
    (`run_images_with_context_augmented`), then the readout of `eval.py:468-476` / `keypoint_regressor.py:196-203`
    as ONE launch over all (image, keypoint) planes.  -> (n, K, 2) (row, col)."""
    maps = run_images_with_context_augmented(
        ldm, [s["img"] for s in samples], context, indices, device=device, from_where=from_where, layers=layers,
        noise_level=noise_level, augmentation_iterations=augmentation_iterations, augment_degrees=augment_degrees,
        augment_scale=augment_scale, augment_translate=augment_translate, controllers=controllers, num_gpus=num_gpus,
        upscale_size=upscale_size, seeds=seeds)
    n, k, h, w = maps.shape
    flat = maps.reshape(n * k, h, w)
    kp = find_max_pixel(flat) if max_loc_strategy == "argmax" else pixel_from_weighted_avg(flat)
    return kp.reshape(n, k, 2) / float(upscale_size)


@torch.no_grad()
def evaluate(
    ldm,
    context,
    indices,
    regressor,
    device="cuda",
    from_where=["down_cross", "mid_cross", "up_cross"],
    upsample_res=32,
    layers=[0, 1, 2, 3, 4, 5],
    noise_level=-1,
    num_tokens=1000,
    augment_degrees=30,
    augment_scale=(0.9, 1.1),
    augment_translate=(0.1, 0.1),
    augmentation_iterations=20,
    dataset_loc="~",
    save_folder="outputs",
    wandb_log=False,
    visualize=False,
    dataset_name="celeba_aligned",
    evaluation_method="inter_eye_distance",
    controllers=None,
    num_gpus=1,
    max_loc_strategy="argmax",
    validation=False,
    dataset=None,
    log_fn=print,
    images_per_pass=1,
    seed=None,
):
    """`eval.py:374-539`: TTA extraction -> readout -> linear regressor -> dataset metric, per test
    image; writes `all_errors.pt` into `save_folder` like the reference and returns the (N,) errors.
    Images are sharded across ranks, one all_gather of the scalar errors at the end.
    `images_per_pass` images share each UNet pass (the reference: 1); `seed` makes image i draw its augmentations and
    noise from generators seeded seed + i, so the result does not depend on the sharding or the grouping."""
    import os
    from . import distributed as dist_
    from .optimize import _make_dataset

    world, rank = dist_.world_size(), dist_.rank()
    dataset = _make_dataset(dataset_name, dataset_loc, dataset, split="test")      # `eval.py:401-424`: the test split
    n = len(dataset)
    per = (n + world - 1) // world
    regressor = torch.as_tensor(regressor).to(device=device, dtype=torch.float32)
    errs = []
    mine = list(dist_.shard_range(n, rank, world))
    for c in range(0, len(mine), max(1, int(images_per_pass))):
        ids = mine[c:c + max(1, int(images_per_pass))]
        samples = [dataset[i] for i in ids]
        kps = extract_keypoints(
            ldm, samples, context, torch.as_tensor(indices).cpu(), device=device, from_where=from_where, layers=layers,
            noise_level=noise_level, augmentation_iterations=augmentation_iterations, augment_degrees=augment_degrees,
            augment_scale=augment_scale, augment_translate=augment_translate, controllers=controllers, num_gpus=num_gpus,
            max_loc_strategy=max_loc_strategy, seeds=None if seed is None else [seed + i for i in ids])
        for sample, kp in zip(samples, kps):
            est = (((kp.reshape(1, -1) - 0.5) @ regressor) + 0.5).reshape(-1, 2)
            gt = torch.as_tensor(sample["kpts"]).to(device=device, dtype=torch.float32)
            vis = torch.as_tensor(sample["visibility"]).to(device) if "visibility" in sample else None
            errs.append(keypoint_error(est, gt, evaluation_method, vis))
    t = torch.stack(errs) if errs else torch.zeros(0, device=device)
    if world > 1:
        t = torch.cat([t, torch.zeros(per - t.shape[0], device=device)])
        t = dist_.interleave_shards(dist_.all_gather_cat(t), n)
    all_errors = t.cpu()
    if rank == 0:
        if log_fn:
            log_fn(f"mean distance: {float(all_errors.mean()) if n else float('nan'):.6f} over {n} images ({evaluation_method})")
        os.makedirs(save_folder, exist_ok=True)
        torch.save(all_errors, os.path.join(save_folder, "all_errors.pt"))
    return all_errors
