"""Stage driver with the reference's on-disk artefacts and resume semantics
(`unsupervised_keypoints/main.py:198-416`).

Stages, in order: optimize -> find_indices -> precompute -> evaluate.  `start_from_stage` names the
first stage to RUN; everything before it is loaded from `save_folder`, written by an earlier run:

    embedding.pt         (1, T, ctx) float32     optimize
    indices.pt           (K,) int64              find_indices
    source_keypoints.pt  (N, K, 2) float32       precompute
    target_keypoints.pt  (N, n, 2) float32       precompute
    visible.pt           (N, n) float32 | None   precompute
    regressor.pt         (2K, 2n) float32        evaluate (fit on the precomputed keypoints)
    all_errors.pt        (N_test,) float32       evaluate

Same file names, shapes and dtypes as the reference, so either implementation can resume from the
other's folder.  Datasets are passed as objects (`train_dataset=` / `test_dataset=`; the reference's
dataset classes are out of scope), visualisation stages are skipped.  Under `torchrun` every rank
runs the stages on its shard; rank 0 writes the files.

    python -m stablekeypoints_b200.main --model_type sd21-base --save_folder out --num_steps 500   # synthetic data
"""
from __future__ import annotations

import argparse
import os
from types import SimpleNamespace

import torch

from . import distributed as dist_
from .eval import evaluate
from .keypoint_regressor import (find_best_indices, precompute_all_keypoints, return_regressor,
                                 return_regressor_human36m, return_regressor_visible)
from .optimize import optimize_embedding
from .optimize_token import load_ldm

STAGES = ("optimize", "find_indices", "precompute", "evaluate")

# the reference's CLI defaults (`main.py:23-196`) for the options the hot path reads
DEFAULTS = dict(
    model_type="runwayml/stable-diffusion-v1-5", dataset_loc="~", save_folder="outputs", dataset_name="celeba_aligned",
    max_len=-1, start_from_stage="optimize", device="cuda:0", lr=5e-3, num_steps=500, num_tokens=500,
    feature_upsample_res=128, batch_size=4, top_k_strategy="gaussian", max_loc_strategy="argmax",
    evaluation_method="inter_eye_distance", min_dist=0.1, furthest_point_num_samples=25, num_indices=100,
    num_subjects=1, sharpening_loss_weight=100.0, equivariance_attn_loss_weight=1000.0, layers=[0, 1, 2, 3],
    noise_level=-1, max_num_points=50_000, sigma=2.0, augment_degrees=15.0, augment_scale=(0.8, 1.0),
    augment_translate=(0.25, 0.25), augmentation_iterations=10, top_k=10, validation=False, wandb=False, visualize=False,
)


def _load(save_folder, name, device):
    obj = torch.load(os.path.join(save_folder, name), map_location="cpu")
    return None if obj is None else obj.to(device).detach()


def _save(obj, save_folder, name):
    if dist_.rank() == 0:
        torch.save(obj if obj is None else obj.detach().cpu(), os.path.join(save_folder, name))
    dist_.barrier()


def fit_regressor(source_kpts, target_kpts, visible, evaluation_method):
    """`main.py:342-366`: which regressor goes with which metric; fp64 fit, float32 result."""
    n = source_kpts.shape[0]
    x = source_kpts.reshape(n, -1).double()
    y = target_kpts.reshape(n, -1).double()
    if evaluation_method in ("visible", "mean_average_error"):
        v = torch.ones_like(y) if visible is None else visible.unsqueeze(-1).repeat(1, 1, 2).reshape(n, -1).double()
        w = return_regressor_visible(x, y, v)
    elif evaluation_method == "orientation_invariant":
        w = return_regressor_human36m(x, y)
    else:
        w = return_regressor(x, y)
    return w.to(torch.float32)


def run_stages(args=None, *, train_dataset=None, test_dataset=None, ldm=None, log_fn=print, **overrides):
    """Runs the stages from `args.start_from_stage` on; returns a dict of every artefact."""
    cfg = dict(DEFAULTS)
    if args is not None:
        cfg.update(vars(args) if not isinstance(args, dict) else args)
    cfg.update(overrides)
    a = SimpleNamespace(**cfg)
    if a.start_from_stage not in STAGES:
        raise ValueError(f"start_from_stage must be one of {STAGES}")
    first = STAGES.index(a.start_from_stage)
    dist_.init_from_env()
    ldm, controllers, num_gpus = load_ldm(a.device, a.model_type, feature_upsample_res=a.feature_upsample_res, ldm=ldm)
    if dist_.rank() == 0:
        os.makedirs(a.save_folder, exist_ok=True)
    dist_.barrier()
    common = dict(device=a.device, layers=a.layers, noise_level=a.noise_level, controllers=controllers, num_gpus=num_gpus)
    aug = dict(augment_degrees=a.augment_degrees, augment_scale=a.augment_scale, augment_translate=a.augment_translate)
    out = {}

    if first <= 0:
        out["embedding"] = optimize_embedding(
            ldm, top_k_strategy=a.top_k_strategy, wandb_log=False, lr=a.lr, num_steps=int(a.num_steps), num_tokens=a.num_tokens,
            top_k=a.top_k, dataset_loc=a.dataset_loc, sigma=a.sigma, sharpening_loss_weight=a.sharpening_loss_weight,
            equivariance_attn_loss_weight=a.equivariance_attn_loss_weight, batch_size=a.batch_size, dataset_name=a.dataset_name,
            max_len=a.max_len, furthest_point_num_samples=a.furthest_point_num_samples, min_dist=a.min_dist,
            validation=a.validation, num_subjects=a.num_subjects, dataset=train_dataset,
            context_dim=getattr(a, "context_dim", None), log_fn=log_fn if log_fn else False, **common, **aug).detach()
        _save(out["embedding"], a.save_folder, "embedding.pt")
    else:
        out["embedding"] = _load(a.save_folder, "embedding.pt", a.device)

    if first <= 1:
        out["indices"] = find_best_indices(
            ldm, out["embedding"], num_steps=a.num_indices, num_tokens=a.num_tokens, top_k=a.top_k, dataset_loc=a.dataset_loc,
            dataset_name=a.dataset_name, min_dist=a.min_dist, top_k_strategy=a.top_k_strategy,
            furthest_point_num_samples=a.furthest_point_num_samples, sigma=a.sigma, validation=a.validation,
            num_subjects=a.num_subjects, dataset=train_dataset, **common)
        _save(out["indices"], a.save_folder, "indices.pt")
    else:
        out["indices"] = _load(a.save_folder, "indices.pt", a.device)

    if first <= 2:
        src, tgt, vis = precompute_all_keypoints(
            ldm, out["embedding"], out["indices"], augmentation_iterations=a.augmentation_iterations, dataset_loc=a.dataset_loc,
            dataset_name=a.dataset_name, max_num_points=a.max_num_points, max_loc_strategy=a.max_loc_strategy,
            save_folder=a.save_folder, validation=a.validation, dataset=train_dataset, **common, **aug)
        out.update(source_keypoints=src, target_keypoints=tgt, visible=vis)
        for name in ("source_keypoints", "target_keypoints", "visible"):
            _save(out[name], a.save_folder, name + ".pt")
    else:
        for name in ("source_keypoints", "target_keypoints", "visible"):
            out[name] = _load(a.save_folder, name + ".pt", a.device)

    out["regressor"] = fit_regressor(out["source_keypoints"], out["target_keypoints"], out["visible"], a.evaluation_method)
    _save(out["regressor"], a.save_folder, "regressor.pt")
    out["all_errors"] = evaluate(
        ldm, out["embedding"], out["indices"], out["regressor"].to(a.device), augmentation_iterations=a.augmentation_iterations,
        dataset_loc=a.dataset_loc, save_folder=a.save_folder, dataset_name=a.dataset_name, evaluation_method=a.evaluation_method,
        max_loc_strategy=a.max_loc_strategy, validation=a.validation, dataset=test_dataset if test_dataset is not None else train_dataset,
        log_fn=log_fn, **common, **aug)
    return out


def _parser():
    p = argparse.ArgumentParser(description="optimize a class embedding (B200 hot path, reference stage layout)")
    for k, v in DEFAULTS.items():
        if isinstance(v, bool):
            p.add_argument(f"--{k}", action="store_true")
        elif isinstance(v, (list, tuple)):
            p.add_argument(f"--{k}", type=type(v[0]), nargs="+", default=list(v))
        else:
            p.add_argument(f"--{k}", type=type(v), default=v)
    p.add_argument("--synthetic_images", type=int, default=64, help="length of the synthetic stand-in dataset")
    return p


if __name__ == "__main__":
    from .datasets import SyntheticKeypoints
    ns = _parser().parse_args()
    ds = SyntheticKeypoints(length=ns.synthetic_images, image_size=512, seed=0)
    run_stages(ns, train_dataset=ds, test_dataset=SyntheticKeypoints(length=max(4, ns.synthetic_images // 8), image_size=512, seed=1))
