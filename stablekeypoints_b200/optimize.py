"""Map collection, losses and the token-optimisation loop — host mirror of the hot-path part
of the reference's `unsupervised_keypoints/optimize.py`, on the kernels of libskp_b200.so.

Reference lines replaced:
  collect_maps                 optimize.py:27-79
  equivariance_loss            optimize.py:157-163
  sharpening_loss              optimize.py:166-179
  find_gaussian_loss_at_point  optimize.py:182-206
  optimize_embedding           optimize.py:269-475
"""
from __future__ import annotations

import time
from typing import Optional, Sequence

import torch

from . import ops
from .invertable_transform import RandomAffineWithInverse


# --------------------------------------------------------------------------------------
# collect_maps (`optimize.py:27-79`)
# --------------------------------------------------------------------------------------
def _selected_handles(controller, layers):
    store = controller.step_store["attn"]
    return [store[i] for i in range(len(store)) if i in layers]


def _layer_args(handles):
    qs = [h.q for h in handles]
    ks = [h.k for h in handles]
    heads = [h.heads for h in handles]
    scales = [h.scale for h in handles]
    return qs, ks, heads, scales, handles[0].feature_res


def collect_maps_batched(controller, from_where=["up_cross"], upsample_res=512, layers=[0, 1, 2, 3], indices=None,
                         precision: str = "auto") -> torch.Tensor:
    """(B, T', R', R') maps for every image of the last UNet pass.

    One fused kernel: bicubic q upsample, Q.K^T, softmax over tokens, mean over heads then
    layers, optional token gather.  The bilinear upsample (`optimize.py:63-70`) is linear, so it
    is applied once to the mean instead of per (layer, head)."""
    handles = _selected_handles(controller, layers)
    if not handles:
        raise RuntimeError("collect_maps: the controller holds no captured layer among `layers`")
    qs, ks, heads, scales, R = _layer_args(handles)
    if indices is not None:
        indices = torch.as_tensor(indices).to(device=qs[0].device, dtype=torch.int64)
    maps = ops.attention_maps(qs, ks, heads, scales, R, token_idx=indices, precision=precision)
    if upsample_res != -1 and upsample_res != R:
        if maps.requires_grad:
            raise RuntimeError("collect_maps: the bilinear upsample is only used by the no-grad stages "
                               "(find_best_indices, evaluation); call under torch.no_grad()")
        maps = ops.bilinear_upsample(maps, int(upsample_res))
    controller.reset()
    return maps


def collect_maps(controller, from_where=["up_cross"], upsample_res=512, layers=[0, 1, 2, 3], indices=None):
    """`optimize.py:27-79` -> (T', R', R').  Same per-image semantics as the reference, which
    holds exactly one image per controller; with a batch in the controller
    use `collect_maps_batched`."""
    maps = collect_maps_batched(controller, from_where, upsample_res, layers, indices)
    if maps.shape[0] != 1:
        raise RuntimeError(f"collect_maps: controller holds {maps.shape[0]} images; use collect_maps_batched "
                           "(the reference's mean over dim 0 would mix images, optimize.py:75)")
    return maps[0]


# --------------------------------------------------------------------------------------
# losses (`optimize.py:157-206`)
# --------------------------------------------------------------------------------------
def _arange_like(n, t):
    return torch.arange(n, device=t.device, dtype=torch.int64)


def equivariance_loss(embeddings_initial, embeddings_transformed, transform, index):
    """MSE(a, un-warp(b)[index]) (`optimize.py:157-163`); gradients into both maps."""
    theta_inv = transform.inverse_theta()[int(index)]
    b = embeddings_transformed[int(index)] if embeddings_transformed.dim() == 4 else embeddings_transformed
    _, eq = ops.keypoint_losses(embeddings_initial, b, _arange_like(embeddings_initial.shape[0], embeddings_initial),
                                theta_inv, 1.0)
    return eq


def find_gaussian_loss_at_point(attn_map, pos, sigma=1.0, temperature=1e-1, device="cuda", indices=None, num_subjects=1):
    """`optimize.py:182-206` with an explicit `pos` (S,N,2) in [0,1]: MSE against gaussian_circles."""
    target = ops.gaussian_circles(pos.to(attn_map.device).float().contiguous(), attn_map.shape[-2], float(sigma))
    if indices is not None:
        attn_map, target = attn_map[indices], target[indices]
    return torch.nn.functional.mse_loss(attn_map, target)


def sharpening_loss(attn_map, sigma=1.0, temperature=1e1, device="cuda", num_subjects=1):
    """`optimize.py:166-179`: MSE(map, gaussian at the map's own argmax).  The argmax search,
    the Gaussian synthesis and the squared error are one kernel; gradient 2(a-g)/(K*H*W)."""
    sharp, _ = ops.keypoint_losses(attn_map, None, _arange_like(attn_map.shape[0], attn_map), None,
                                   float(sigma), int(num_subjects))
    return sharp


# --------------------------------------------------------------------------------------
# token optimisation (`optimize.py:269-475`)
# --------------------------------------------------------------------------------------
# --------------------------------------------------------------------------------------
# helpers the reference defines but never calls from `main` (`optimize.py:82-135, 209-266`;
# SURVEY.md §8(a) row 19).  Kept for API completeness; plain differentiable tensor algebra on
# the caller's device, not part of the measured path.
# --------------------------------------------------------------------------------------
def create_gaussian_kernel(size: int, sigma: float) -> torch.Tensor:
    """Normalised (size, size) Gaussian, centre at size // 2 (`optimize.py:82-103`)."""
    assert size % 2 == 1, "Size must be odd"
    r = torch.arange(size, dtype=torch.float32) - size // 2
    k = torch.exp(-(r[:, None] ** 2 + r[None, :] ** 2) / (2 * sigma ** 2))
    return k / k.sum()


def gaussian_loss(attn_map, kernel_size=5, sigma=1.0, temperature=1e-4):
    """`optimize.py:106-135`: MSE between the map and its own temperature pixel-softmax (gradient
    stopped).  The reference also blurs that softmax with a 5x5 Gaussian and max-normalises it, then
    does not use the result (`:120-132`); the blur is therefore not computed here."""
    t = attn_map.shape[0]
    probs = torch.softmax(attn_map.reshape(t, -1) / temperature, dim=1).detach().reshape(attn_map.shape)
    return torch.nn.functional.mse_loss(attn_map, probs)


def _normalised(heatmaps):
    return heatmaps / (heatmaps.sum(dim=(1, 2), keepdim=True) + 1e-6)


def variance_loss(heatmaps):
    """Mean over planes of the spatial standard deviation of the normalised heat-map (`optimize.py:209-241`)."""
    _, m, n = heatmaps.shape
    p = _normalised(heatmaps)
    x = torch.arange(m, device=heatmaps.device, dtype=torch.float32).view(1, m, 1)
    y = torch.arange(n, device=heatmaps.device, dtype=torch.float32).view(1, 1, n)
    xm = (x * p).sum(dim=(1, 2), keepdim=True)
    ym = (y * p).sum(dim=(1, 2), keepdim=True)
    return torch.sqrt((p * ((x - xm) ** 2 + (y - ym) ** 2)).sum(dim=(1, 2))).mean()


def differentiable_argmax(heatmaps):
    """Soft-argmax `sum(coord * v) / (sum(v) + 1e-6)` -> (N, 2) as (row, col), no half-pixel offset
    (`optimize.py:244-266`).  The readout used by the live path is `eval.pixel_from_weighted_avg`."""
    _, m, n = heatmaps.shape
    p = _normalised(heatmaps)
    x = torch.arange(m, device=heatmaps.device, dtype=torch.float32).view(1, m, 1)
    y = torch.arange(n, device=heatmaps.device, dtype=torch.float32).view(1, 1, n)
    return torch.stack([(x * p).sum(dim=(1, 2)), (y * p).sum(dim=(1, 2))], dim=1)


def select_tokens(attn_map, attn_map_transformed, top_k_strategy, furthest_point_num_samples, top_k, sigma, num_subjects):
    """Candidate ranking on the image's maps + furthest-point sampling on the transformed image's
    maps (`optimize.py:405-418`) -> (top_k,) int64 on the maps' device, no host sync."""
    from . import ptp_utils
    if top_k_strategy == "entropy":
        cand = ptp_utils.entropy_sort(attn_map, furthest_point_num_samples)
    elif top_k_strategy == "gaussian":
        cand = ptp_utils.find_top_k_gaussian(attn_map, furthest_point_num_samples, sigma=sigma, num_subjects=num_subjects)
    elif top_k_strategy == "consistent":
        cand = torch.arange(furthest_point_num_samples, device=attn_map.device)
    else:
        raise NotImplementedError
    return ptp_utils.furthest_point_sampling(attn_map_transformed, top_k, cand)


def select_tokens_batched(maps, maps_transformed, top_k_strategy, furthest_point_num_samples, top_k, sigma, num_subjects):
    """`select_tokens` for the B images of a batch with one launch per stage (statistics over all B*T planes, one
    ranking CTA and one furthest-point CTA per image).  maps / maps_transformed (B, T, H, W) -> (B, top_k) int64."""
    B, T, H, W = maps.shape
    n_cand = min(int(furthest_point_num_samples), T)
    flat = maps.reshape(B * T, H, W)
    if top_k_strategy == "entropy":
        _, score, _ = ops.plane_stats(flat, want_kl=False, want_entropy=True)
    elif top_k_strategy == "gaussian":
        score, _, _ = ops.plane_stats(flat, sigma=float(sigma), epsilon=1e-5, num_subjects=int(num_subjects), want_kl=True)
    elif top_k_strategy == "consistent":
        score = None
    else:
        raise NotImplementedError
    cand = (torch.arange(n_cand, device=maps.device).expand(B, -1).contiguous() if score is None
            else ops.rank_ascending_batched(score.reshape(B, T), n_cand))
    am = ops.plane_argmax(maps_transformed.reshape(B * T, H, W)).reshape(B, T)
    return ops.furthest_point_sampling_batched(am, H, W, cand, int(top_k))


def token_opt_losses(ldm, image, transformed_img, context, transform, controllers, *, layers, noise_level, from_where,
                     device, top_k_strategy, furthest_point_num_samples, top_k, sigma, num_subjects, noise=None,
                     generator=None, precision="auto"):
    """One image-iteration of `optimize.py:372-429` for a batch of B images on this device.

    Both UNet passes of the reference (image, transformed image) run as ONE pass over 2B images
    (every normalisation in the UNet is per-sample, so batching does not change any value).
    Selection and losses of all B images are one launch per stage; nothing is read back to the host.
    `precision` selects the arithmetic of the attention maps and of their backward (`ops.attention_maps`): "auto" = tcgen05
    (forward fp32-grade bf16x3; gradient products on single-bf16 operands, <= 1.5e-2 of each gradient's max, measured <= 7e-3);
    "simt_f32" = the fp32 CUDA-core kernels in both directions (2e-3; ~80 x slower: the A/B of
    `tests/test_e2e_gpu.py::test_optimisation_converges_alike_with_tensor_core_and_fp32_backward`).
    -> (mean sharpening loss, mean equivariance loss, idx (B, top_k))."""
    from . import ptp_utils
    B = image.shape[0]
    ctrl = ptp_utils._controller_list(controllers)[0]
    ctrl.reset()
    both = torch.cat([image, transformed_img], dim=0)
    ptp_utils.find_pred_noise(ldm, both, context, noise_level=noise_level, device=device, noise=noise, generator=generator)
    handles = _selected_handles(ctrl, layers)
    qs, ks, heads, scales, R = _layer_args(handles)
    ctrl.reset()
    with torch.no_grad():
        maps, stats, fwd_ws = ops.attention_maps([q.detach() for q in qs], [k.detach() for k in ks], heads, scales, R,
                                                 precision=precision, return_stats=True, return_workspace=True)   # (2B,T,R,R)
        idx = select_tokens_batched(maps[:B], maps[B:], top_k_strategy, furthest_point_num_samples, top_k, sigma, num_subjects)
    # the backward reuses the native-resolution logits / interpolation tables the forward staged in its workspace
    sel = ops.selected_planes(maps, torch.cat([idx, idx], dim=0), qs, ks, heads, scales, R, precision=precision,
                              row_stats=stats, fwd_workspace=fwd_ws)                                      # (2B,K,R,R)
    per_image = ops.keypoint_losses_batched(sel, transform.inverse_theta(), float(sigma), int(num_subjects))
    return per_image[:, 0].mean(), per_image[:, 1].mean(), idx


def _make_dataset(dataset_name, dataset_loc, dataset=None, split="train", max_len=-1, **kw):
    """`dataset=` wins; else the name is resolved like the reference's stages do (`optimize.py:301-326`, `eval.py:401-424`)."""
    if dataset is not None:
        return dataset
    from . import datasets as D
    if dataset_name == "synthetic":
        return D.SyntheticKeypoints(**kw)
    return D.make_dataset(dataset_name, dataset_loc, split=split, max_len=max_len)


def optimize_embedding(
    ldm,
    top_k_strategy="entropy",
    wandb_log=True,
    context=None,
    device="cuda",
    num_steps=2000,
    from_where=["down_cross", "mid_cross", "up_cross"],
    upsample_res=256,
    layers=[0, 1, 2, 3, 4, 5],
    lr=5e-3,
    noise_level=-1,
    num_tokens=1000,
    top_k=10,
    augment_degrees=30,
    augment_scale=(0.9, 1.1),
    augment_translate=(0.1, 0.1),
    dataset_loc="~",
    sigma=1.0,
    sharpening_loss_weight=100,
    equivariance_attn_loss_weight=100,
    batch_size=4,
    num_gpus=1,
    dataset_name="celeba_aligned",
    max_len=-1,
    min_dist=0.05,
    furthest_point_num_samples=50,
    controllers=None,
    validation=False,
    num_subjects=1,
    dataset=None,
    context_dim=None,
    log_fn=None,
    step_hook=None,
    seed=None,
    grad_hook=None,
    precision="auto",
):
    """`optimize.py:269-475` with the same arguments (+ `precision`, see `token_opt_losses`).  `num_gpus` keeps the reference's meaning
    "images per UNet pass" (its DataLoader batch, `optimize.py:356`); here those images share
    ONE GPU.  Across processes (torchrun, one per GPU) every rank runs this function on its own
    shard of each batch and the context gradient is all-reduced (mean) before each optimiser
    step (`distributed.allreduce_mean_`), replacing `nn.DataParallel` (`optimize_token.py:43-44`).
    Random draws (augmentation, noise, batch order) come from per-rank generators seeded `seed + rank`
    (`distributed.rank_generators`; `seed=None` in a single process keeps the global generators).
    `grad_hook(context.grad)` is called after the all-reduce, before the optimiser step.
    Returns the optimised context, detached (identical on every rank)."""
    from . import distributed as dist_
    from . import ptp_utils
    world, rank = dist_.world_size(), dist_.rank()
    dataset = _make_dataset(dataset_name, dataset_loc, dataset, split="train", max_len=max_len)
    transform = RandomAffineWithInverse(degrees=augment_degrees, scale=augment_scale, translate=augment_translate)
    if context is None:
        dim = context_dim or getattr(getattr(ldm, "cfg", None), "context_dim", 768)
        context = ptp_utils.init_random_noise(device, num_words=num_tokens, dim=dim)
    context = context.detach().to(device).clone()
    dist_.broadcast_(context)
    context.requires_grad = True
    optimizer = torch.optim.Adam([context], lr=lr)
    if wandb_log:
        import wandb   # noqa: F401  (same dependency as the reference; off by default in our drivers)
    per_pass = num_gpus
    if batch_size % (per_pass * world) != 0:
        raise ValueError(f"batch_size {batch_size} must be a multiple of num_gpus*world_size = {per_pass * world}")
    accum = batch_size // (per_pass * world)
    host_data = not torch.as_tensor(dataset[0]["img"]).is_cuda
    g_cpu, g_dev = dist_.rank_generators(device, seed)
    shard = dist_.shard_dataset(dataset, rank, world)
    if len(shard) < per_pass:
        raise ValueError(f"rank {rank}: {len(shard)} image(s) of the {len(dataset)}-image dataset, fewer than the {per_pass} of one UNet "
                         f"pass (num_gpus): the dataset must hold at least num_gpus * world_size = {per_pass * world} images")
    loader = torch.utils.data.DataLoader(shard, batch_size=per_pass, shuffle=True, drop_last=True, pin_memory=host_data,
                                         generator=g_cpu)
    it = iter(loader)
    start = it_start = time.time()
    run_eq = run_sh = run_tot = 0.0
    for iteration in range(int(num_steps * accum)):
        try:
            mini_batch = next(it)
        except StopIteration:
            it = iter(loader)
            mini_batch = next(it)
        image = mini_batch["img"].to(device, non_blocking=True)
        transformed = transform(image, generator=g_cpu)
        sh, eq, _ = token_opt_losses(
            ldm, image, transformed, context, transform, controllers, layers=layers, noise_level=noise_level,
            from_where=from_where, device=device, top_k_strategy=top_k_strategy,
            furthest_point_num_samples=furthest_point_num_samples, top_k=top_k, sigma=sigma, num_subjects=num_subjects,
            generator=g_dev, precision=precision)
        loss = eq * equivariance_attn_loss_weight + sh * sharpening_loss_weight
        run_eq = run_eq + eq.detach() / accum * equivariance_attn_loss_weight
        run_sh = run_sh + sh.detach() / accum * sharpening_loss_weight
        run_tot = run_tot + loss.detach() / accum
        (loss / accum).backward()
        if (iteration + 1) % accum == 0:
            dist_.allreduce_mean_(context.grad)
            if grad_hook is not None:
                grad_hook(context.grad)
            optimizer.step()
            optimizer.zero_grad()
            if log_fn is False:
                pass                      # silent: no device->host read, the step stays asynchronous
            elif log_fn is not None or not wandb_log:
                rec = {"loss": float(run_tot), "running_equivariance_attn_loss": float(run_eq),
                       "running_sharpening_loss": float(run_sh), "iteration time": time.time() - it_start}
                if log_fn is not None:
                    log_fn(rec)
                elif rank == 0:
                    print(rec)
            elif rank == 0:
                import wandb
                wandb.log({"loss": float(run_tot), "running_equivariance_attn_loss": float(run_eq),
                           "running_sharpening_loss": float(run_sh), "iteration time": time.time() - it_start})
            run_eq = run_sh = run_tot = 0.0
            it_start = time.time()
            if step_hook is not None:
                step_hook((iteration + 1) // accum)
    if rank == 0 and log_fn is not False:
        print(f"optimization took {time.time() - start} seconds")
    return context.detach()
