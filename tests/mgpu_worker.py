"""torchrun worker of tests/test_multi_gpu.py (one process per GPU, NCCL).  Not collected by pytest.

Checks, on hardware, the multi-GPU semantics that replace the reference's nn.DataParallel (`optimize.py:428-429`:
the per-GPU losses are averaged, so the step's gradient is the mean of the per-image gradients):
  1. the all-reduced context gradient of one token-optimisation step at world N == the single-process gradient of
     the same (image, theta, noise) triples: (a) accumulated over passes of the same composition as the ranks' shards
     (identical kernels on identical shapes: only atomics order differs) to 1e-5 of the max, and (b) on the
     concatenated batch in ONE pass to 1e-3 of the max (the stock trunk picks other cuBLAS / cuDNN kernels for another
     batch size, and those fp32 differences pass through the bf16 gradient products);
  2. `precompute_all_keypoints` (image-sharded, no data-path collective) returns bit-identical keypoints at world N
     and in a single process;
  3. per-rank random streams differ (augmentation thetas, noise), and `optimize_embedding` leaves every rank with the
     same context.
Rank 0 writes {"ok": bool, ...} as JSON to argv[1].
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(out_path):
    from stablekeypoints_b200 import distributed as D
    from stablekeypoints_b200 import keypoint_regressor, optimize, optimize_token
    from stablekeypoints_b200.datasets import SyntheticKeypoints
    from stablekeypoints_b200.invertable_transform import RandomAffineWithInverse
    from stablekeypoints_b200.sd_standin import PipelineStandIn, unet_config

    rank, local_rank, world = D.init_from_env("nccl")
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    cfg = unet_config("tiny-sd21")
    pipe = PipelineStandIn(cfg, vae_channels=(32, 32, 64, 64), seed=0).to(dev)
    ldm, controllers, _ = optimize_token.load_ldm(dev, feature_upsample_res=32, ldm=pipe)
    T, per_rank = 48, 2
    n_img = per_rank * world
    g = torch.Generator().manual_seed(7)
    context0 = torch.randn(1, T, cfg.context_dim, generator=g)
    ds = SyntheticKeypoints(length=n_img, image_size=128, seed=5)
    images = torch.stack([ds[i]["img"] for i in range(n_img)])
    tf = RandomAffineWithInverse(degrees=15, scale=(0.8, 1.0), translate=(0.25, 0.25))
    thetas = tf.sample_theta(n_img, generator=g)
    lat = cfg.latent_side
    noise = torch.randn(2 * n_img, 4, lat, lat, generator=g)       # [image i | transformed copy n_img + i]
    kw = dict(layers=[0, 1, 2, 3], noise_level=-1, from_where=None, top_k_strategy="gaussian",
              furthest_point_num_samples=12, top_k=5, sigma=2.0, num_subjects=1)

    def step_gradient(ids, allreduce):
        ctx = context0.to(dev).clone().requires_grad_(True)
        img = images[ids].to(dev)
        transformed = tf(img, theta=thetas[ids])
        nz = torch.cat([noise[ids], noise[[n_img + i for i in ids]]]).to(dev)
        sh, eq, idx = optimize.token_opt_losses(ldm, img, transformed, ctx, tf, controllers, device=dev, noise=nz, **kw)
        (eq * 1000.0 + sh * 100.0).backward()
        if allreduce:
            D.allreduce_mean_(ctx.grad)
        return ctx.grad.detach().clone(), idx

    mine = list(D.shard_range(n_img))
    g_dist, idx_dist = step_gradient(mine, True)
    g_single, idx_single = step_gradient(list(range(n_img)), False)          # every rank computes the references itself
    g_accum = sum(step_gradient(list(D.shard_range(n_img, r, world)), False)[0] for r in range(world)) / world
    scale = float(g_single.abs().max())
    err = float((g_dist - g_accum).abs().max()) / scale
    err_cat = float((g_dist - g_single).abs().max()) / scale
    idx_ok = bool(torch.equal(idx_dist.cpu(), idx_single.cpu()[mine]))

    # 2. extraction: sharded vs unsharded, seeded per image
    top = torch.arange(5) * 9
    ds5 = SyntheticKeypoints(length=2 * world + 1, image_size=128, seed=6)     # n % world != 0: uneven shards
    kwe = dict(device=dev, layers=[0, 1, 2, 3], augmentation_iterations=3, augment_degrees=15, augment_scale=(0.8, 1.0),
               augment_translate=(0.25, 0.25), controllers=controllers, num_gpus=3, dataset=ds5, seed=1234)
    # same images per UNet pass in both runs: identical kernels on identical shapes, so the keypoints must be bit-identical
    src_d, tgt_d, vis_d = keypoint_regressor.precompute_all_keypoints(ldm, context0.to(dev), top, images_per_pass=1, **kwe)
    # n < world: a rank with an empty shard still enters every gather with the right shapes (no hang, no mismatch)
    src_1, tgt_1, vis_1 = keypoint_regressor.precompute_all_keypoints(ldm, context0.to(dev), top, max_num_points=1, **kwe)
    # single-process reference: bypass the process group for the same call
    real_ws, real_rank = D.world_size, D.rank
    D.world_size, D.rank = (lambda: 1), (lambda: 0)
    try:
        src_s, tgt_s, vis_s = keypoint_regressor.precompute_all_keypoints(ldm, context0.to(dev), top, images_per_pass=1, **kwe)
    finally:
        D.world_size, D.rank = real_ws, real_rank
    kp_equal = (bool(torch.equal(src_d.cpu(), src_s.cpu())) and bool(torch.equal(tgt_d.cpu(), tgt_s.cpu()))
                and bool(torch.equal(vis_d.cpu(), vis_s.cpu())) and src_d.shape == (2 * world + 1, 5, 2)
                and bool(torch.equal(src_1.cpu(), src_s[:1].cpu())) and tgt_1.shape == (1, 5, 2) and vis_1.shape == (1, 5))

    # 3. per-rank random streams + identical context after optimize_embedding
    g_cpu, g_dev = D.rank_generators(dev, seed=99)
    my_theta = tf.sample_theta(1, g_cpu).to(dev)
    my_noise = torch.randn(4, device=dev, generator=g_dev)
    all_theta = D.all_gather_cat(my_theta)
    all_noise = D.all_gather_cat(my_noise[None])
    streams_differ = all(not torch.equal(all_theta[0], all_theta[r]) and not torch.equal(all_noise[0], all_noise[r])
                         for r in range(1, world))
    grads = []
    ctx_out = optimize.optimize_embedding(
        ldm, top_k_strategy="gaussian", wandb_log=False, context=context0, device=dev, num_steps=2, layers=[0, 1, 2, 3], lr=5e-3,
        num_tokens=T, top_k=5, augment_degrees=15, augment_scale=(0.8, 1.0), augment_translate=(0.25, 0.25), sigma=2.0,
        sharpening_loss_weight=100, equivariance_attn_loss_weight=1000, batch_size=2 * world, num_gpus=2,
        furthest_point_num_samples=12, controllers=controllers, dataset=ds, log_fn=False, seed=99,
        grad_hook=lambda gr: grads.append(gr.detach().clone()))
    all_ctx = D.all_gather_cat(ctx_out)
    all_g = D.all_gather_cat(grads[0])
    ctx_same = all(torch.equal(all_ctx[0], all_ctx[r]) for r in range(1, world))
    grad_same = all(torch.equal(all_g[0], all_g[r]) for r in range(1, world))
    moved = float((ctx_out - context0.to(dev)).abs().max()) > 0

    res = torch.tensor([err, err_cat, float(idx_ok), float(kp_equal), float(streams_differ), float(ctx_same), float(grad_same), float(moved)],
                       device=dev, dtype=torch.float64)
    allres = D.all_gather_cat(res[None])
    if rank == 0:
        rec = {"world": world, "grad_err_vs_same_passes": float(allres[:, 0].max()),
               "grad_err_vs_concatenated_batch": float(allres[:, 1].max()), "grad_scale": scale,
               "idx_ok": bool(allres[:, 2].min()), "keypoints_identical": bool(allres[:, 3].min()),
               "streams_differ": bool(allres[:, 4].min()), "context_identical": bool(allres[:, 5].min()),
               "allreduced_grad_identical": bool(allres[:, 6].min()), "context_moved": bool(allres[:, 7].min())}
        rec["ok"] = (rec["grad_err_vs_same_passes"] <= 1e-5 and rec["grad_err_vs_concatenated_batch"] <= 1e-3
                     and rec["idx_ok"] and rec["keypoints_identical"]
                     and rec["streams_differ"] and rec["context_identical"] and rec["allreduced_grad_identical"]
                     and rec["context_moved"])
        with open(out_path, "w") as f:
            json.dump(rec, f)
        print(json.dumps(rec), flush=True)
    D.barrier()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1])
