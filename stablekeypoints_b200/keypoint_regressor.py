"""Token voting and keypoint precomputation — host mirror of the hot-path part of the
reference's `unsupervised_keypoints/keypoint_regressor.py`.

Reference lines replaced:
  find_best_indices         keypoint_regressor.py:16-121
  precompute_all_keypoints  keypoint_regressor.py:124-224
"""
from __future__ import annotations

import torch

from . import distributed as dist_
from . import ops
from .eval import extract_keypoints
from .optimize import _make_dataset, select_tokens_batched


def vote_top_k(indices_list: torch.Tensor, top_k: int) -> torch.Tensor:
    """`keypoint_regressor.py:114-119`: most frequently selected token ids.  The reference's
    `counts.argsort(descending=True)` leaves the order of equally-voted ids unspecified; here
    ties go to the smaller id (stable sort), which is one of the orders it may produce."""
    ids, counts = torch.unique(indices_list.reshape(-1).cpu(), return_counts=True)
    order = torch.sort(counts, descending=True, stable=True).indices
    return ids[order][:top_k]


@torch.no_grad()
def find_best_indices(
    ldm,
    context,
    num_steps=100,
    device="cuda",
    noise_level=-1,
    upsample_res=256,
    layers=[0, 1, 2, 3, 4, 5],
    from_where=["down_cross", "mid_cross", "up_cross"],
    num_tokens=1000,
    top_k=30,
    dataset_loc="~",
    dataset_name="celeba_aligned",
    min_dist=0.05,
    furthest_point_num_samples=50,
    controllers=None,
    num_gpus=1,
    top_k_strategy="entropy",
    sigma=3,
    validation=False,
    num_subjects=1,
    dataset=None,
):
    """`keypoint_regressor.py:16-121` -> (top_k,) int64 CPU tensor.  Images are sharded across
    ranks (no collective on the data path); the per-image votes are gathered once."""
    from . import ptp_utils
    world, rank = dist_.world_size(), dist_.rank()
    dataset = _make_dataset(dataset_name, dataset_loc, dataset)
    loader = torch.utils.data.DataLoader(dist_.shard_dataset(dataset, rank, world), batch_size=num_gpus,
                                         shuffle=True, drop_last=True)
    it = iter(loader)
    votes = []
    for _ in range(num_steps // (num_gpus * world)):
        try:
            mini_batch = next(it)
        except StopIteration:
            it = iter(loader)
            mini_batch = next(it)
        image = mini_batch["img"].to(device, non_blocking=True)
        maps = ptp_utils.run_and_find_attn_batched(
            ldm, image, context, layers=layers, noise_level=noise_level, from_where=from_where,
            upsample_res=upsample_res, device=device, controllers=controllers)
        votes.append(select_tokens_batched(maps, maps, top_k_strategy, furthest_point_num_samples, top_k, sigma, num_subjects))
    votes = torch.cat(votes) if votes else torch.empty(0, top_k, dtype=torch.int64, device=device)
    votes = dist_.all_gather_cat(votes)
    return vote_top_k(votes, top_k)


@torch.no_grad()
def precompute_all_keypoints(
    ldm,
    context,
    top_indices,
    device="cuda",
    noise_level=-1,
    layers=[0, 1, 2, 3, 4, 5],
    from_where=["down_cross", "mid_cross", "up_cross"],
    augment_degrees=30,
    augment_scale=(0.9, 1.1),
    augment_translate=(0.1, 0.1),
    augmentation_iterations=20,
    dataset_loc="~",
    visualize=False,
    dataset_name="celeba_aligned",
    controllers=None,
    num_gpus=1,
    max_num_points=50_000,
    max_loc_strategy="argmax",
    save_folder="outputs",
    validation=False,
    dataset=None,
    images_per_pass=1,
    seed=None,
):
    """`keypoint_regressor.py:124-224` -> (source (N,K,2), target (N,n,2), visibility|None), in
    dataset order.  Images are sharded across ranks; one all_gather of the (N_local,K,2) keypoints
    at the end (the reference shuffles with batch_size 1; order does not affect the regressor).
    `images_per_pass` images share each UNet pass (the reference: 1); `seed`: image i draws its augmentations and noise
    from generators seeded seed + i (results independent of sharding / grouping)."""
    world, rank = dist_.world_size(), dist_.rank()
    dataset = _make_dataset(dataset_name, dataset_loc, dataset)
    n = min(len(dataset), max_num_points)
    mine = list(dist_.shard_range(n, rank, world))
    per = (n + world - 1) // world
    src, tgt, vis = [], [], []
    # every rank issues the same collectives with equal shapes, whatever its shard holds: the shapes come from sample 0
    # and the argument list, not from the local rows (a rank with an empty shard must neither skip a gather nor guess)
    first = dataset[0] if n > 0 else {}
    has_vis = "visibility" in first
    n_kpts = tuple(torch.as_tensor(first["kpts"]).shape) if n > 0 else (0, 2)
    step = max(1, int(images_per_pass))
    for c in range(0, len(mine), step):
        ids = mine[c:c + step]
        samples = [dataset[i] for i in ids]
        kps = extract_keypoints(
            ldm, samples, context, top_indices, device=device, from_where=from_where, layers=layers, noise_level=noise_level,
            augmentation_iterations=augmentation_iterations, augment_degrees=augment_degrees, augment_scale=augment_scale,
            augment_translate=augment_translate, controllers=controllers, num_gpus=num_gpus, max_loc_strategy=max_loc_strategy,
            seeds=None if seed is None else [seed + i for i in ids])
        for sample, kp in zip(samples, kps):
            src.append(kp)
            tgt.append(torch.as_tensor(sample["kpts"]).to(device=device, dtype=torch.float32))
            if has_vis:
                vis.append(torch.as_tensor(sample["visibility"]).to(device=device, dtype=torch.float32))

    def _gather(rows, row_shape):
        t = torch.stack(rows) if rows else torch.zeros(0, *row_shape, device=device)
        if world == 1:
            return t if n > 0 else None
        if t.shape[0] < per:                    # pad the short shard so all_gather shapes agree
            t = torch.cat([t, torch.zeros(per - t.shape[0], *row_shape, device=device, dtype=t.dtype)])
        return dist_.interleave_shards(dist_.all_gather_cat(t), n)

    return (_gather(src, (len(top_indices), 2)), _gather(tgt, n_kpts),
            _gather(vis, n_kpts[:1]) if has_vis else None)


# --------------------------------------------------------------------------------------
# linear regressors from unsupervised to annotated keypoints (`keypoint_regressor.py:227-299`)
# Closed-form least squares in fp64 on the tensors' own device; (N, 2K) -> (N, 2n), both
# centred at 0.5 like the reference.  Accept numpy arrays or tensors, return what they were given.
# --------------------------------------------------------------------------------------
def _as64(a):
    was_numpy = not torch.is_tensor(a)
    return torch.as_tensor(a).to(torch.float64), was_numpy


def _back(w: torch.Tensor, was_numpy: bool):
    return w.cpu().numpy() if was_numpy else w


def _lstsq_pinv(x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    return torch.linalg.pinv(x.T @ x) @ x.T @ y


def return_regressor(X, Y):
    """`W = pinv(X^T X) X^T Y` on X - 0.5, Y - 0.5 (`keypoint_regressor.py:256-266`)."""
    x, np_in = _as64(X)
    y, _ = _as64(Y)
    return _back(_lstsq_pinv(x - 0.5, y.to(x.device) - 0.5), np_in)


def return_regressor_visible(X, Y, visible):
    """One least-squares problem per output column over the rows where that column is visible
    (`keypoint_regressor.py:227-253`)."""
    x, np_in = _as64(X)
    y, _ = _as64(Y)
    v, _ = _as64(visible)
    x, y, v = x - 0.5, y.to(x.device) - 0.5, v.to(x.device)
    w = torch.zeros(x.shape[1], y.shape[1], dtype=torch.float64, device=x.device)
    for j in range(y.shape[1]):
        rows = torch.nonzero(v[:, j] == 1).flatten()
        w[:, j] = _lstsq_pinv(x[rows], y[rows, j])
    return _back(w, np_in)


def return_regressor_human36m(X, Y):
    """Least squares with the Human3.6M left/right label swap: refit until at most 10 samples are
    closer to their mirrored annotation than to the given one (`keypoint_regressor.py:269-299`)."""
    from .eval import swap_points
    x, np_in = _as64(X)
    y, _ = _as64(Y)
    x, y = x - 0.5, (y.to(x.device) - 0.5).clone()
    proj = torch.linalg.inv(x.T @ x) @ x.T
    while True:
        w = proj @ y
        pred = x @ w
        n = x.shape[0]
        dist = (pred - y).reshape(n, -1, 2).norm(dim=2).mean(dim=1)
        swapped = swap_points(y.reshape(n, -1, 2)).reshape(n, -1)
        swapped_dist = (pred - swapped).reshape(n, -1, 2).norm(dim=2).mean(dim=1)
        should_swap = dist > swapped_dist
        if int(should_swap.sum()) > 10:
            y[should_swap] = swapped[should_swap]
        else:
            break
    return _back(w, np_in)
